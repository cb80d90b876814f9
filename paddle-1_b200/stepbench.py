"""Single-GPU step pipeline shared by bench.py (N = 1) and the sharded bench's same-batch baseline.

One step = lookup_table forward of a batch, then the fused sort/duplicate-reduction/optimizer update with the
SelectedRows gradient aliasing (ids, dOut) — b200emb_lookup_table_f32 + b200emb_sparse_sgd_f32 /
b200emb_sparse_adam_f32 through the C ABI.  The sort plan of batch i+1 needs the ids only (a data-loader
prefetch): it runs on side streams beside gather/update of batch i, every step still launching exactly one
plan, one gather and one update.  Steps are replayed as CUDA graphs over `n_rotate` batch buffers.
"""
import torch

from . import capi


class SingleGpuPipeline:
    def __init__(self, table, ids, dout, optimizer="sgd", lr=0.01, status=None, prefetch=True):
        """table [V, D] float32 on the GPU (updated in place); ids / dout: lists of device tensors, one per
        rotating batch buffer (together they should exceed L2)."""
        self.table, self.ids, self.dout = table, ids, dout
        self.nb = len(ids)
        self.dev = table.device
        self.v, self.d = table.shape
        self.n = ids[0].numel()
        self.opt = optimizer
        self.prefetch = prefetch
        self.lr = torch.tensor([lr], dtype=torch.float32, device=self.dev)
        self.status = status if status is not None else capi.new_status(self.dev)
        self.out = [torch.empty((self.n, self.d), dtype=torch.float32, device=self.dev) for _ in range(self.nb)]
        self.ws = [capi.alloc_workspace(self.n, self.d, 4, self.dev) for _ in range(self.nb)]
        if optimizer == "adam":
            self.m1, self.m2 = torch.zeros_like(table), torch.zeros_like(table)
            self.b1p = torch.tensor([0.9], device=self.dev)
            self.b2p = torch.tensor([0.999], device=self.dev)
        # the critical path (gather -> reduce/update) gets the high-priority stream; the sort plans of the
        # following batches fill in behind it on normal-priority streams
        self.main = torch.cuda.Stream(self.dev, priority=-1)
        self.sides = [torch.cuda.Stream(self.dev, priority=0) for _ in range(2)]
        self.graphs, self.rotation = None, None

    # ---- the three C-ABI calls of a step
    def plan(self, b, stream=None):
        capi.plan_rows(self.ids[b], self.v, self.ws[b], self.status, stream=stream)

    def gather(self, b):
        capi.lookup_table(self.table, self.ids[b], -1, self.out[b], self.status)

    def update(self, b, flags):
        if self.opt == "adam":
            capi.sparse_adam(self.table, self.m1, self.m2, self.ids[b], self.dout[b], self.lr, self.b1p, self.b2p,
                             workspace=self.ws[b], flags=flags, status=self.status)
        else:
            capi.sparse_sgd(self.table, self.ids[b], self.dout[b], self.lr, self.ws[b], flags, self.status)

    def kernels_per_step(self):
        # gather, sort plan, chunk reduce (the fast-mode update finds the run heads itself: no head scan)
        return 1 + capi.plan_kernel_count(self.n, self.v) + 1

    def _step(self, b):
        """One step on batch b outside a rotation: plan of batch b+1 on a side stream during this step."""
        nb = (b + 1) % self.nb
        side = self.sides[0]
        if self.prefetch:
            side.wait_stream(self.main)
            with torch.cuda.stream(side):
                self.plan(nb)
            self.gather(b)
            self.update(b, capi.PLAN_READY)
            self.main.wait_stream(side)
        else:
            self.gather(b)
            self.update(b, 0)

    def capture(self):
        main, nb = self.main, self.nb
        with torch.cuda.stream(main):
            for b in range(nb):   # every workspace holds its batch's plan before the first step
                self.plan(b)
            for b in range(nb):   # warm every path once outside capture
                self._step(b)
            main.synchronize()
            self.graphs = []
            for b in range(nb):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=main):
                    self._step(b)
                self.graphs.append(g)
            # one graph holding a whole rotation (nb consecutive steps): amortises the per-graph launch gap.
            # Inside it the plan of batch b+1 only has to be ready when update(b+1) starts, so the plans run as
            # two alternating side chains that overlap gather AND update of the main stream.
            self.rotation = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.rotation, stream=main):
                if self.prefetch:
                    for st in self.sides:
                        st.wait_stream(main)
                    plan_done, upd0_done = {}, None
                    for b in range(nb):
                        nxt = (b + 1) % nb
                        st = self.sides[nxt % 2]
                        with torch.cuda.stream(st):
                            if nxt == 0:
                                st.wait_event(upd0_done)      # ws[0] is still read by this rotation's update(0)
                            self.plan(nxt)
                            plan_done[nxt] = torch.cuda.Event()
                            plan_done[nxt].record(st)
                        self.gather(b)
                        if b in plan_done:
                            main.wait_event(plan_done[b])
                        self.update(b, capi.PLAN_READY)
                        if b == 0:
                            upd0_done = torch.cuda.Event()
                            upd0_done.record(main)
                    for st in self.sides:
                        main.wait_stream(st)
                else:
                    for b in range(nb):
                        self.gather(b)
                        self.update(b, 0)
        torch.cuda.synchronize(self.dev)

    def run_steps(self, k, start=0):
        """k consecutive steps starting at batch `start % nb`: whole rotations as one graph launch, the ragged
        head/tail as single-step graphs.  Call inside `with torch.cuda.stream(self.main)`."""
        i = 0
        while i < k and (start + i) % self.nb != 0:
            self.graphs[(start + i) % self.nb].replay()
            i += 1
        while k - i >= self.nb:
            self.rotation.replay()
            i += self.nb
        while i < k:
            self.graphs[(start + i) % self.nb].replay()
            i += 1

    def time_steps(self, steps, warmup):
        """(ms for `steps` steps, device-timed on the main stream)."""
        with torch.cuda.stream(self.main):
            self.run_steps(warmup)
            torch.cuda.synchronize(self.dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.main)
            self.run_steps(steps, warmup)
            e1.record(self.main)
            torch.cuda.synchronize(self.dev)
        return e0.elapsed_time(e1)

    def time_phase(self, fn, reps):
        """Seconds per call of fn(b), replayed as one CUDA graph per batch buffer on the main stream."""
        main, nb = self.main, self.nb
        with torch.cuda.stream(main):
            for b in range(nb):
                fn(b)
            main.synchronize()
            gs = []
            for b in range(nb):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=main):
                    fn(b)
                gs.append(g)
            for b in range(nb):
                gs[b].replay()
            main.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(main)
            for i in range(reps):
                gs[i % nb].replay()
            a1.record(main)
            main.synchronize()
        return a0.elapsed_time(a1) / reps * 1e-3

    def phase_times(self, reps):
        """{gather, sort_plan, update (PLAN_READY), serial (gather + sort + update back to back)} in seconds."""
        t_gather = self.time_phase(self.gather, reps)
        t_plan = self.time_phase(lambda b: self.plan(b), reps)
        for b in range(self.nb):
            self.plan(b)
        t_update = self.time_phase(lambda b: self.update(b, capi.PLAN_READY), reps)
        t_serial = self.time_phase(lambda b: (self.gather(b), self.update(b, 0)), reps)
        return {"gather": t_gather, "sort_plan": t_plan, "update": t_update, "serial": t_serial}


def hashed_rows(global_rows, d, dtype=torch.float32):
    """Table initialiser both sides can evaluate: value[r, c] = ((r*1315423911 + c*2654435761) mod 1048573)
    / 1048573 * 0.1 - 0.05, computed in int64 / float64 and rounded to float32 — identical bits in torch (any
    device) and numpy, so a checker can regenerate the rows it needs of a table too big to ship."""
    r = global_rows.to(torch.int64).reshape(-1, 1)
    c = torch.arange(d, dtype=torch.int64, device=r.device).reshape(1, -1)
    h = (r * 1315423911 + c * 2654435761) % 1048573
    return (h.to(torch.float64) / 1048573.0 * 0.1 - 0.05).to(dtype)


def init_shard_hashed(shard, rank, world, height, chunk=1 << 21):
    """shard[i, :] = hashed_rows(i*world + rank) for the rows that exist; the padding rows stay zero."""
    hs, d = shard.shape
    for lo in range(0, hs, chunk):
        hi = min(hs, lo + chunk)
        gid = torch.arange(lo, hi, dtype=torch.int64, device=shard.device) * world + rank
        vals = hashed_rows(gid, d, shard.dtype)
        vals[gid >= height] = 0
        shard[lo:hi] = vals
