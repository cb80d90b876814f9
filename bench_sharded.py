"""bench.py --gpus N (N > 1): BASELINE.json configs[3] — 100M x 64 fp32 table row-sharded over the N GPUs of one
box (owner = id % N), Zipf(1.1) ids, sparse SGD.  Launched by torchrun, one rank per GPU.  Weak scaling: every
GPU brings its own batch — by default the SAME per-GPU batch as `bench.py --gpus 1` (4096 x 26 = 106 496 ids),
so the driver's 1 -> N curve compares like with like; `sweep` adds larger per-GPU batches, each beside the
single-GPU step time on the same batch and the same (unsharded) 100M-row table.

The step goes through the C ABI's row-sharded engine (b200emb_shard_*): no collective on the data path.
After the timed runs a checked epilogue replays 8 consecutive steps (prefetch pipeline on) in the fast and in
the bit-exact mode and compares them with the oracle run on the concatenated id streams of all ranks; the
line carries the result in `parity` and the process exits non-zero if it fails.  (bench code: the oracle is
used here as the checker only, never on the timed path.)  Last, BASELINE.json configs[4] — 50M x 256 table,
65536 x 100 duplicate-heavy ids per GPU — runs row-sharded over the same GPUs as a `configs_extra` entry with
its own checked step (c5_sharded); B200EMB_BENCH_NO_C5=1 or B200EMB_BENCH_QUICK=1 skip it.
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

METRIC = "embedding lookups+updates/sec"
UNIT = "ids/s"
N_ROTATE = 4         # batch buffers = engine slots
NVLINK_GBS = 770.0   # measured peer-copy bandwidth per direction per GPU (B200_PROFILING.md)
SWEEP = (4096 * 26 * 4, 4096 * 26 * 16)
PARITY_IDS = 32768
PARITY_STEPS = 8


def _peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class Workload:
    """ids / dOut batches of one per-GPU batch size, resident on the device, plus pinned host copies."""

    def __init__(self, pkg, zipf, n, d, rank, dev, pinned):
        synth = pkg.synth
        self.n = n
        ids_np = [zipf.sample(n) for _ in range(N_ROTATE)]
        self.ids_np = ids_np
        self.ids_h = [torch.from_numpy(i).pin_memory() if pinned else torch.from_numpy(i) for i in ids_np]
        self.ids = [t.to(dev) for t in self.ids_h]
        g = torch.Generator(device=dev)
        g.manual_seed(4321 + rank)
        self.dout = [torch.randn((n, d), device=dev, generator=g) * 0.01 for _ in range(N_ROTATE)]
        self.dout_h = [t.cpu().pin_memory() for t in self.dout] if pinned else None
        self.out = [torch.empty((n, d), dtype=torch.float32, device=dev) for _ in range(N_ROTATE)]


def recv_capacity(ids_np_list, world, rank):
    """Unique rows this owner receives per step (max over the batches), from the host copies of every rank's ids."""
    mine = np.zeros((len(ids_np_list), world), np.int64)
    for b, ids in enumerate(ids_np_list):
        u = np.unique(ids)
        mine[b] = np.bincount(u % world, minlength=world)
    t = torch.from_numpy(mine).cuda()
    allc = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(allc, t)
    tot = torch.stack(allc).sum(dim=0)          # [batch, owner]
    return int(tot.max().item()), int(mine.sum(axis=1).max())   # the same number on every rank (engine layouts must match)


class Stepper:
    """The sharded step with the ids-only part of batch i+1 issued on a side stream before step i is queued,
    eagerly or as CUDA graphs (one per batch buffer for prepare, one for lookup + update)."""

    def __init__(self, tab, wl, lr, dev, use_graphs):
        self.tab, self.wl, self.lr, self.dev = tab, wl, lr, dev
        self.main = torch.cuda.current_stream(dev)
        self.side = torch.cuda.Stream(dev)
        self.graphs = None
        if use_graphs:
            try:
                self._capture()
            except Exception as e:          # graph capture is an optimisation: fall back to eager launches
                print("rank %d: CUDA graph capture failed (%s); eager launches" % (tab.rank, e), file=sys.stderr)
                self.graphs = None
                torch.cuda.synchronize(dev)

    def _capture(self):
        eng, wl = self.tab.engine, self.wl
        for b in range(N_ROTATE):           # warm up every call outside capture
            eng.prepare(b, wl.ids[b])
            eng.lookup(b, wl.out[b])
            eng.update(b, wl.dout[b], self.lr)
        torch.cuda.synchronize(self.dev)
        dist.barrier()
        gp, gs = [], []
        cap = torch.cuda.Stream(self.dev)
        for b in range(N_ROTATE):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                eng.prepare(b, wl.ids[b])
            gp.append(g)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                eng.lookup(b, wl.out[b])
                eng.update(b, wl.dout[b], self.lr)
            gs.append(g)
        torch.cuda.synchronize(self.dev)
        self.graphs = (gp, gs)
        self.ev_prep = [torch.cuda.Event() for _ in range(N_ROTATE)]

    def run(self, k):
        tab, wl = self.tab, self.wl
        if self.graphs is None:
            prep = tab.prepare(wl.ids[0], None, stream=self.side, after=self.main)
            out = None
            for i in range(k):
                b = i % N_ROTATE
                nxt = tab.prepare(wl.ids[(b + 1) % N_ROTATE], None, stream=self.side, after=self.main) if i + 1 < k else None
                out, ctx = tab.lookup(wl.ids[b], None, prep=prep)
                tab.update(ctx, wl.dout[b], self.lr)
                prep = nxt
            return out
        gp, gs = self.graphs

        def prep(b):
            self.side.wait_stream(self.main)
            with torch.cuda.stream(self.side):
                gp[b].replay()
                self.ev_prep[b].record(self.side)
        prep(0)
        for i in range(k):
            b = i % N_ROTATE
            if i + 1 < k:
                prep((b + 1) % N_ROTATE)
            self.main.wait_event(self.ev_prep[b])
            gs[b].replay()
        return wl.out[(k - 1) % N_ROTATE]


def measure(pkg, tab, zipf, n, d, rank, world, dev, steps, warm, lr, want_e2e, use_graphs):
    capi = pkg.capi
    wl = Workload(pkg, zipf, n, d, rank, dev, pinned=want_e2e)
    cap_recv, u_max = recv_capacity(wl.ids_np, world, rank)
    tab.enable_engine(n, int(cap_recv * 1.25) + 1024)
    st = Stepper(tab, wl, lr, dev, use_graphs)
    st.run(warm)
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    st.run(steps)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    tab.check_status()
    bounds, nu, seg, _ = tab.engine.debug_views((steps - 1) % N_ROTATE if st.graphs is not None else tab._slot)
    u_local, recv_rows = int(nu.item()), int(seg[-1].item())
    own = int((bounds[rank + 1] - bounds[rank]).item())
    res = {"n": n, "ms_per_step": ms / steps, "value": world * n * steps / (ms * 1e-3), "unique_local": u_local,
           "recv_rows": recv_rows, "graphs": st.graphs is not None,
           # bytes this rank sends over NVLink per step: ids + gradient rows of the rows it does not own, and the
           # rows it returns to the other requesters (about as many as it receives from them)
           "nvlink_out_bytes": (u_local - own) * (8 + 4 * d) + (recv_rows - own) * 4 * d}
    # per-kernel-group device times on rank 0: eager launches, one engine phase per call, CUDA events around each
    # (consumer phases include the wait for the peers' flags, so rank skew shows up there)
    res["phases_us"] = phase_breakdown(tab.engine, wl, lr, dev, 6)
    dist.barrier()
    if want_e2e:
        res["e2e"] = e2e_pipeline(tab, wl, lr, dev, world, max(8, min(steps, 40)))
    tab.engine.destroy()
    tab.peer = False
    del wl, st
    torch.cuda.empty_cache()
    return res


def phase_breakdown(eng, wl, lr, dev, steps):
    names = ["prepare: sort plan + de-duplication + ids to owners (side stream, one batch ahead)",
             "owner gather -> requesters, rows flags (waits for every source's ids)",
             "fork owner sort plan (side stream)", "expand to batch order (waits for every owner)",
             "local MergeAdd -> owners, gradient flags",
             "owner fused update (waits for every source, joins owner sort plan)"]
    main = torch.cuda.current_stream(dev)
    side = torch.cuda.Stream(dev)
    acc = [0.0] * len(names)
    evs = []
    dist.barrier()
    for i in range(steps + 2):
        b = i % N_ROTATE
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
        side.wait_stream(main)
        with torch.cuda.stream(side):
            ev[0].record(side)
            eng.prepare(b, wl.ids[b])
            ev[1].record(side)
        main.wait_stream(side)
        ev[2].record(main)
        for ph in range(1, 4):
            eng.lookup(b, wl.out[b], phases=(ph, ph))
            ev[2 + ph].record(main)
        for ph in range(2):
            eng.update(b, wl.dout[b], lr, phases=(ph, ph))
            ev[6 + ph].record(main)
        evs.append(ev)
    torch.cuda.synchronize(dev)
    for ev in evs[2:]:
        acc[0] += ev[0].elapsed_time(ev[1])
        for k in range(5):
            acc[1 + k] += ev[2 + k].elapsed_time(ev[3 + k])
    return {n_: round(a / steps * 1e3, 1) for n_, a in zip(names, acc)}


def e2e_pipeline(tab, wl, lr, dev, world, steps):
    """Host batches in, rows out, every step: pinned ids + dOut -> H2D, prepare + lookup + update, Out -> D2H, on
    three streams with double-buffered staging so the copies of neighbouring steps overlap the kernels."""
    n, d = wl.n, wl.out[0].shape[1]
    nbuf = 2
    ids_d = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(nbuf)]
    dout_d = [torch.empty((n, d), dtype=torch.float32, device=dev) for _ in range(nbuf)]
    out_h = [torch.empty((n, d), dtype=torch.float32).pin_memory() for _ in range(nbuf)]
    outs = [None] * nbuf
    s_in, s_cmp, s_out = (torch.cuda.Stream(dev) for _ in range(3))
    ev_in = [torch.cuda.Event() for _ in range(nbuf)]
    ev_cmp = [torch.cuda.Event() for _ in range(nbuf)]
    ev_out = [torch.cuda.Event() for _ in range(nbuf)]

    def one(i):
        b, src = i % nbuf, i % N_ROTATE
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_cmp[b])          # staging buffers are free once the step that used them computed
            ids_d[b].copy_(wl.ids_h[src], non_blocking=True)
            dout_d[b].copy_(wl.dout_h[src], non_blocking=True)
            ev_in[b].record(s_in)
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(ev_in[b])
            s_cmp.wait_event(ev_out[b])         # the previous user of this output buffer has been fetched
            prep = tab.prepare(ids_d[b], None)
            outs[b], ctx = tab.lookup(ids_d[b], None, prep=prep)
            tab.update(ctx, dout_d[b], lr)
            ev_cmp[b].record(s_cmp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_cmp[b])
            out_h[b].copy_(outs[b], non_blocking=True)
            ev_out[b].record(s_out)

    for i in range(4):
        one(i)
    torch.cuda.synchronize(dev)
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    torch.cuda.synchronize(dev)
    el = torch.tensor([time.perf_counter() - t0], device=dev)
    dist.all_reduce(el, op=dist.ReduceOp.MAX)
    el = float(el.item())
    return {"value": world * n * steps / el, "unit": UNIT, "h2d_bytes_per_step": world * (n * 8 + n * d * 4),
            "d2h_bytes_per_step": world * n * d * 4, "ms_per_step": el / steps * 1e3, "steps": steps,
            "api": "ShardedTable.prepare/lookup/update (b200emb_shard_* C ABI) on pinned host batches, 3-stream "
                   "feed/fetch pipeline per GPU"}


def one_gpu_same_batch(pkg, zipf, v, d, n, dev, steps):
    """Rank 0 only: the single-GPU step (bench.py's pipeline) on the same per-GPU batch and the unsharded table."""
    from importlib import import_module
    stepbench = import_module("paddle1_b200.stepbench")
    table = torch.empty((v, d), dtype=torch.float32, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    for lo in range(0, v, 1 << 22):
        table[lo:lo + (1 << 22)].uniform_(-0.05, 0.05, generator=g)
    nb = 8 if n <= 131072 else 3
    ids = [torch.from_numpy(zipf.sample(n)).to(dev) for _ in range(nb)]
    dout = [torch.randn((n, d), device=dev, generator=g) * 0.01 for _ in range(nb)]
    pipe = stepbench.SingleGpuPipeline(table, ids, dout, "sgd", 0.01)
    pipe.capture()
    ms = pipe.time_steps(steps, max(10, 2 * nb))
    del pipe, table, ids, dout
    torch.cuda.empty_cache()
    return {"ms_per_step": ms / steps, "value": n * steps / (ms * 1e-3)}


# --------------------------------------------------------------------------- checked epilogue
def parity_epilogue(pkg, tab, zipf_seeded, v, d, rank, world, dev):
    """PARITY_STEPS consecutive steps of PARITY_IDS ids per rank, prefetch pipeline on, in the fast mode (engine)
    and in the bit-exact mode (collective path), each from a freshly initialised shard, against the oracle on
    the concatenated id streams: the last step's gathered rows, every touched row this rank owns (incl. its
    hottest one).  Every rank regenerates every rank's inputs from the seeds, so nothing big crosses ranks."""
    import oracle
    from importlib import import_module
    stepbench = import_module("paddle1_b200.stepbench")
    capi, synth, sharded = pkg.capi, pkg.synth, pkg.sharded
    n, steps = PARITY_IDS, PARITY_STEPS
    lr_v = np.array([0.05], np.float32)
    lr = torch.tensor([0.05], device=dev)
    ids_all = [[zipf_seeded(1000 + r * 64 + s).sample(n) for r in range(world)] for s in range(steps)]
    g_all = [[synth.grad_values(n, d, seed=9000 + r * 64 + s) for r in range(world)] for s in range(steps)]
    # the oracle works on the touched rows only (the table has 100M rows): compact index via np.unique
    touched = np.unique(np.concatenate([np.concatenate(x) for x in ids_all]))
    init = stepbench.hashed_rows(torch.from_numpy(touched), d).numpy()
    want_det = init.copy()
    truth = init.astype(np.float64)
    bound = np.zeros_like(truth)
    want_last_out = last_truth = last_bound = None
    for s in range(steps):
        ci = [np.searchsorted(touched, i) for i in ids_all[s]]
        if s == steps - 1:
            want_last_out = want_det[ci[rank]].copy()
            last_truth, last_bound = truth[ci[rank]].copy(), bound[ci[rank]].copy()
        cat_i, cat_g = np.concatenate(ci), np.concatenate(g_all[s])
        oracle.sgd_sparse(want_det, cat_i, cat_g, lr_v)         # reference order: device-major concat, input order
        rows, merged = oracle.merge_add(cat_i, cat_g, truth=True)
        _, l1 = oracle.merge_add(cat_i, np.abs(cat_g), truth=True)
        truth[rows] -= 0.05 * merged
        bound[rows] += 1e-6 * 0.05 * l1
    mine = touched % world == rank
    local_rows = torch.from_numpy(touched[mine] // world).to(dev)
    ids_d = [torch.from_numpy(ids_all[s][rank]).to(dev) for s in range(steps)]
    g_d = [torch.from_numpy(g_all[s][rank]).to(dev) for s in range(steps)]
    side = torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    res = {"steps": steps, "ids_per_rank_per_step": n, "touched_rows_checked": int(mine.sum())}

    def run(det, ws):
        stepbench.init_shard_hashed(tab.shard, rank, world, v)
        torch.cuda.synchronize(dev)
        dist.barrier()
        prep = tab.prepare(ids_d[0], ws[0], det, stream=side, after=main)
        out = None
        for s in range(steps):
            nxt = tab.prepare(ids_d[s + 1], ws[(s + 1) % 4], det, stream=side, after=main) if s + 1 < steps else None
            out, ctx = tab.lookup(ids_d[s], ws[s % 4], deterministic=det, prep=prep)
            tab.update(ctx, g_d[s], lr, deterministic=det)
            prep = nxt
        torch.cuda.synchronize(dev)
        dist.barrier()
        return out.cpu().numpy(), tab.shard[local_rows].cpu().numpy()

    # fast mode through the engine
    cap_recv = max(sum(int((np.unique(i) % world == o).sum()) for i in ids_all[s]) for s in range(steps) for o in range(world))
    tab.enable_engine(n, int(cap_recv * 1.25) + 1024)
    out, rows_after = run(False, [None] * 4)
    tab.check_status()
    tab.engine.destroy()
    tab.peer = False
    # forward of the last step saw 7 updates: inside the accumulated reordering bound of the fp64 evaluation
    err_f = np.abs(rows_after - truth[mine])
    ok_rows = bool((err_f <= bound[mine] + 1e-6 * np.abs(truth[mine]) + 1e-30).all())
    ref32 = want_det[mine]
    rel = np.abs(rows_after - ref32) / np.maximum(np.abs(ref32), 1e-30)
    res.update({"fast_rows_within_bound": ok_rows, "fast_max_rel_vs_ref_fp32": float(rel.max()),
                "fast_frac_gt_1e-6_rel_vs_ref_fp32": float((rel > 1e-6).mean()),
                "fast_last_out_max_abs_err": float(np.abs(out - want_last_out).max())})
    # the last step's gathered rows saw steps-1 fast-mode updates: inside the accumulated bound of the fp64 evaluation
    ok_out = bool((np.abs(out - last_truth) <= last_bound + 1e-6 * np.abs(last_truth) + 1e-30).all())
    res["fast_last_out_within_bound"] = ok_out
    ok_fast = ok_rows and ok_out
    # bit-exact mode through the collective path (every occurrence shipped, owners apply them in reference order)
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(4)]
    out, rows_after = run(True, ws)
    det_rows = bool(np.array_equal(rows_after.view(np.uint32), want_det[mine].view(np.uint32)))
    det_out = bool(np.array_equal(out.view(np.uint32), want_last_out.view(np.uint32)))
    flags = torch.tensor([int(ok_fast), int(det_rows and det_out)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    mx = torch.tensor([res["fast_max_rel_vs_ref_fp32"], res["fast_frac_gt_1e-6_rel_vs_ref_fp32"],
                       res["fast_last_out_max_abs_err"]], device=dev, dtype=torch.float64)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    res.update({"fast_ok": bool(flags[0].item()), "det_bit_exact": bool(flags[1].item()),
                "fast_max_rel": float(mx[0].item()), "fast_frac_gt_1e-6_rel_vs_ref_fp32": float(mx[1].item()),
                "fast_last_out_max_abs_err": float(mx[2].item()),
                "note": "max over ranks; fast mode = engine path, bit-exact mode = collective path; reference = "
                        "oracle (CPU restatement of the reference kernels) on the concatenated id streams"})
    res.pop("fast_max_rel_vs_ref_fp32", None)
    return res


# --------------------------------------------------------------------------- configs[4] at N GPUs
def c5_sharded(pkg, rank, world, dev, peak, steps):
    """BASELINE.json configs[4] on N GPUs: the duplicate-heavy batch (65536 x 100 Zipf(1.3) ids per GPU per step, one
    id with ~1.67 M duplicates) on the 50M x 256 fp32 table row-sharded over the GPUs, sparse SGD, through the same
    engine.  A timed run, then one checked step from a hashed table: the gathered rows must be the table's bits,
    and sampled updated rows (the hottest id among them) must sit inside the reordering bound of the fp64 sum of
    every rank's gradient rows for that id (partial sums per rank, all-reduced: nothing big crosses ranks)."""
    from importlib import import_module
    stepbench = import_module("paddle1_b200.stepbench")
    capi, synth, sharded = pkg.capi, pkg.synth, pkg.sharded
    v, d, n, a, opt = synth.CONFIGS["C5_dup_heavy"]
    tab = sharded.ShardedTable(v, d, sharded.CudaOps(capi), dev)
    stepbench.init_shard_hashed(tab.shard, rank, world, v)
    lr = torch.tensor([0.01], dtype=torch.float32, device=dev)
    zipf = synth.ZipfIds(v, a, seed=100 + rank)
    use_graphs = os.environ.get("B200EMB_SHARD_NO_GRAPH") is None
    r = measure(pkg, tab, zipf, n, d, rank, world, dev, steps, 4, lr, False, use_graphs)
    # ---- one checked step
    stepbench.init_shard_hashed(tab.shard, rank, world, v)
    ids_np = zipf.sample(n)
    ids = torch.from_numpy(ids_np).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(777 + rank)
    dout = torch.randn((n, d), device=dev, generator=g) * 0.01
    n_s = 4097
    if rank == 0:
        u, c = np.unique(ids_np, return_counts=True)
        pick = np.random.default_rng(5).choice(u, size=n_s - 1, replace=False)
        samp = np.unique(np.concatenate([[u[np.argmax(c)]], pick]))
        samp = np.concatenate([samp, np.full(n_s - len(samp), -1, np.int64)])
        hot = int(c.max())
    else:
        samp, hot = np.zeros(n_s, np.int64), 0
    st = torch.from_numpy(samp).to(dev)
    dist.broadcast(st, 0)
    samp_d = st[st >= 0]
    idx = torch.searchsorted(samp_d, ids).clamp_(max=samp_d.numel() - 1)
    hit = (samp_d[idx] == ids).nonzero().squeeze(1)
    part = torch.zeros((samp_d.numel(), d), dtype=torch.float64, device=dev)
    l1 = torch.zeros_like(part)
    for lo in range(0, hit.numel(), 1 << 19):       # bounded temporaries: the hottest id alone has > 1 M matches
        h = hit[lo:lo + (1 << 19)]
        gd = dout[h].double()
        part.index_add_(0, idx[h], gd)
        l1.index_add_(0, idx[h], gd.abs())
    dist.all_reduce(part)
    dist.all_reduce(l1)
    cap_recv, _ = recv_capacity([ids_np], world, rank)
    tab.enable_engine(n, int(cap_recv * 1.25) + 1024)
    torch.cuda.synchronize(dev)
    dist.barrier()
    prep = tab.prepare(ids, None)
    out, ctx = tab.lookup(ids, None, prep=prep)
    tab.update(ctx, dout, lr)
    torch.cuda.synchronize(dev)
    dist.barrier()
    tab.check_status()
    chk = torch.randint(0, n, (65536,), device=dev)
    gather_ok = bool(torch.equal(out[chk].view(torch.int32), stepbench.hashed_rows(ids[chk], d).view(torch.int32)))
    mine = samp_d % world == rank
    init = stepbench.hashed_rows(samp_d[mine], d).double()
    truth = init - 0.01 * part[mine]
    got = tab.shard[samp_d[mine] // world].double()
    bound = 1e-6 * (0.01 * l1[mine] + init.abs()) + 1e-30
    rows_ok = bool(((got - truth).abs() <= bound).all().item())
    moved = bool(((got - init).abs().amax(dim=1) > 0).all().item()) if int(mine.sum().item()) else True
    flags = torch.tensor([int(gather_ok), int(rows_ok), int(moved)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    tab.engine.destroy()
    tab.peer = False
    ms = r["ms_per_step"]
    b_req = n * (8 + 4 * d) + n * (8 + 4 * d)
    b_own = r["recv_rows"] * (8 + 2 * 4 * d) + r["recv_rows"] * (8 + 4 * d) + r["recv_rows"] * 2 * 4 * d
    hbm = (b_req + b_own) / (ms * 1e-3) / 1e9
    nv = r["nvlink_out_bytes"] / (ms * 1e-3) / 1e9
    entry = {"config": "C5_dup_heavy, row-sharded over %d GPUs" % world, "table": [v, d], "ids_per_gpu_per_step": n,
             "optimizer": opt, "ms_per_step": ms, "value": r["value"], "unit": UNIT, "steps": steps,
             "unique_ids_per_gpu": r["unique_local"], "rows_received_per_gpu": r["recv_rows"],
             "cuda_graphs": r["graphs"], "phases_us": r["phases_us"],
             "roofline": {"hbm_GBs_per_gpu": hbm, "hbm_frac": hbm / peak, "nvlink_GBs_per_direction": nv,
                          "nvlink_frac": nv / NVLINK_GBS},
             "parity": {"checked": "one step from a hashed table: 65536 gathered rows bit-exact; %d sampled updated rows "
                                   "(incl. the hottest id, %d duplicates on rank 0 alone) within 1e-6 * (lr * sum|g| + |p|) "
                                   "of the fp64 sum over every rank's gradient rows; bit-exact mode not run at this size"
                                   % (int(samp_d.numel()), hot),
                        "gather_bit_exact": bool(flags[0].item()), "fast_rows_within_bound": bool(flags[1].item()),
                        "sampled_rows_updated": bool(flags[2].item()),
                        "ok": bool(flags.min().item())}}
    del tab, ids, dout, out, part, l1
    torch.cuda.empty_cache()
    return entry


def run(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus:
        raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    pkg = load_package()
    capi, synth, sharded = pkg.capi, pkg.synth, pkg.sharded
    capi.lib()
    peak, peak_src = _peak()
    v, d, n_base, a, opt = synth.CONFIGS["C4_sharded_sgd"]
    n_main = int(getattr(args, "ids_per_gpu", 0) or n_base)
    ops = sharded.CudaOps(capi)
    tab = sharded.ShardedTable(v, d, ops, dev)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    for lo in range(0, tab.hs, 1 << 22):
        tab.shard[lo:min(tab.hs, lo + (1 << 22))].uniform_(-0.05, 0.05, generator=g)
    lr = torch.tensor([0.01], dtype=torch.float32, device=dev)
    zipf = synth.ZipfIds(v, a, seed=rank)
    use_graphs = os.environ.get("B200EMB_SHARD_NO_GRAPH") is None
    warm = max(args.warmup, 10)
    main = measure(pkg, tab, zipf, n_main, d, rank, world, dev, args.steps, warm, lr, True, use_graphs)
    base = {}
    sweep = []
    quick = os.environ.get("B200EMB_BENCH_QUICK") is not None
    for n in ([] if quick else [x for x in SWEEP if x != n_main]):
        r = measure(pkg, tab, zipf, n, d, rank, world, dev, max(10, min(args.steps, 40)), 10, lr, False, use_graphs)
        sweep.append(r)
    if not quick:
        for n in [n_main] + [r["n"] for r in sweep]:
            if rank == 0:
                base[n] = one_gpu_same_batch(pkg, zipf, v, d, n, dev, max(20, min(args.steps, 100 if n <= 131072 else 30)))
            dist.barrier()

    def zipf_seeded(seed):
        z = synth.ZipfIds.__new__(synth.ZipfIds)
        z.height, z.a, z.cdf, z.mult = zipf.height, zipf.a, zipf.cdf, zipf.mult
        z.rng = np.random.Generator(np.random.PCG64(seed))
        return z
    parity = parity_epilogue(pkg, tab, zipf_seeded, v, d, rank, world, dev)
    ok = parity["fast_ok"] and parity["det_bit_exact"]
    hs_rows = tab.hs
    extra = []
    if not quick and os.environ.get("B200EMB_BENCH_NO_C5") is None:
        del tab.shard                  # configs[4] needs the room: 51 GB / N per GPU plus 4 x 13.4 GB of batch buffers
        torch.cuda.empty_cache()
        c5 = c5_sharded(pkg, rank, world, dev, peak, max(4, min(args.steps, 8)))
        extra.append(c5)
        ok = ok and c5["parity"]["ok"]

    if rank == 0:
        def roof(r):
            n, ms = r["n"], r["ms_per_step"]
            # per-GPU algorithmic HBM bytes: requester side (ids, expand write, dOut read) + owner side (gather
            # read+write of the received rows, their gradient rows, read+write of the rows updated)
            b_req = n * (8 + 4 * d) + n * (8 + 4 * d)
            b_own = r["recv_rows"] * (8 + 2 * 4 * d) + r["recv_rows"] * (8 + 4 * d) + r["recv_rows"] * 2 * 4 * d
            hbm = (b_req + b_own) / (ms * 1e-3) / 1e9
            nv = r["nvlink_out_bytes"] / (ms * 1e-3) / 1e9
            return {"hbm_GBs_per_gpu": hbm, "hbm_frac": hbm / peak, "nvlink_GBs_per_direction": nv,
                    "nvlink_frac": nv / NVLINK_GBS}
        rf = roof(main)
        plan_k = capi.plan_kernel_count(n_main, tab.plan_h)
        # prepare: plan, head scan, inverse map, shard split, send ids | lookup: owner gather, owner plan, expand |
        # update: local MergeAdd, owner update
        launches = (plan_k + capi.HEAD_SCAN_KERNELS + 3) + (1 + capi.plan_kernel_count(main["recv_rows"], tab.hs) + 1) + 2

        def entry(r):
            e = {"ids_per_gpu_per_step": r["n"], "ms_per_step": r["ms_per_step"], "value": r["value"], "unit": UNIT,
                 "unique_ids_per_gpu": r["unique_local"], "rows_received_per_gpu": r["recv_rows"],
                 "cuda_graphs": r["graphs"], "roofline": roof(r), "phases_us": r["phases_us"]}
            if r["n"] in base:
                e["one_gpu_same_batch"] = base[r["n"]]
                e["speedup_vs_one_gpu_same_batch"] = r["value"] / base[r["n"]]["value"]
            return e
        line = {
            "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": main["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[3]: synthetic 100Mx64 fp32 table row-sharded over %d B200 (owner = id %% G), "
                                   "%d Zipf(1.1) ids per GPU per step (the per-GPU batch of configs[1]), ids / rows / "
                                   "gradients exchanged over NVLink by the kernels themselves, sparse SGD" % (world, n_main),
                       "table": [v, d], "ids_per_gpu_per_step": n_main, "unique_ids_per_gpu": main["unique_local"],
                       "rows_received_per_gpu": main["recv_rows"], "optimizer": opt,
                       "mode": "fast (fixed-tree reduction), duplicates merged before they cross NVLink; row-sharded "
                               "engine b200emb_shard_*: peer stores + per-source arrival flags, no collective, no host read",
                       "l2": "inputs rotate over %d batch buffers per GPU (%.0f MB) > 126 MB L2 for the sweep sizes; "
                             "the shard is %.1f GB" % (N_ROTATE, N_ROTATE * n_main * (8 + 8 * d) / 1e6, hs_rows * d * 4 / 1e9),
                       "launch": ("CUDA graphs per batch buffer (prepare | lookup+update); " if main["graphs"] else "eager launches; ") +
                                 "the ids-only part of step i+1 (sort plan, de-duplication) runs on a side stream during "
                                 "step i (data-loader prefetch); every step still does one of each"},
            "roofline": {"bound": "hbm", "achieved": rf["hbm_GBs_per_gpu"], "peak": peak, "unit": "GB/s",
                         "frac": rf["hbm_frac"], "traffic": None, "peak_source": peak_src,
                         "kernel": "whole sharded step, per GPU (requester + owner side)",
                         "nvlink": {"achieved_GBs_per_direction": rf["nvlink_GBs_per_direction"], "peak": NVLINK_GBS,
                                    "frac": rf["nvlink_frac"]}},
            "e2e": main["e2e"], "phases_us": main["phases_us"], "gpu_launches": launches * args.steps,
            "one_gpu_same_batch": base.get(n_main), "sweep": [entry(main)] + [entry(r) for r in sweep],
            "parity": parity, "configs_extra": extra,
        }
        if n_main in base:
            line["speedup_vs_one_gpu_same_batch"] = main["value"] / base[n_main]["value"]
        print(json.dumps(line))
        sys.stdout.flush()
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        raise SystemExit("sharded parity FAILED: %s %s" % (json.dumps(parity), json.dumps([e["parity"] for e in extra])))
