"""Row-sharded embedding table over the GPUs of one box.

Replaces ParallelExecutor's multi-device SelectedRows path — gather every device's sparse gradient
onto one device (`framework/details/reduce_op_handle.cc:88-94`, `reduce_and_gather.h:56-93`), update,
broadcast the parameter (`broadcast_op_handle.cc:71-122`) with the table replicated everywhere — by
the sharding convention of the reference's own distributed lookup table: owner(id) = id % G, local
row = id // G (`operators/split_ids_op.h:45-49`), rows returned in the requester's order
(`operators/merge_ids_op.h:63-87`).

One process per GPU.  Per step and rank:
  forward   plan+dedup local ids by (owner, local row)  ->  unique local rows to their owners (all-to-all)
            -> owners gather rows from their shard -> unique rows back (all-to-all) -> expand to batch order
  backward  fast mode: local MergeAdd on the same plan (one row per unique id, already in send order)
                       -> merged rows to owners (all-to-all) -> owners run the fused sparse SGD/Adam
            bit-exact mode: no pre-merge; every (row id, gradient row) goes to its owner in
                       (owner, row, position) order, owners apply them in (source rank, position) order,
                       which is the reference's gather-concat order.
Two implementations of the step live here:
  * the ENGINE path (default on one box, fast mode): `enable_engine()` hands the whole step to the C ABI's
    b200emb_shard_* (csrc/shard_engine.cu) — kernels that store straight into the other GPUs' buffers and meet
    through arrival flags; no collective, no host read, nothing allocated per step;
  * the COLLECTIVE path below (bit-exact mode, gloo/CPU tests, B200EMB_NO_PEER=1): torch.distributed
    all-to-alls around the same kernels; its only host synchronisation per step is the read of the split sizes.

`ops` is the compute backend (CudaOps = the C-ABI kernels; tests inject a CPU checker backend to
exercise this file's logic with gloo), `torch.distributed` is the communication backend.
"""
import contextlib
import os

import torch
import torch.distributed as dist


def shard_height(height, world):
    return (int(height) + world - 1) // world


class CudaOps:
    """The product backend: every call is a kernel of libb200emb.so (no fallback)."""

    def __init__(self, capi):
        self.capi = capi
        capi.lib()

    def alloc_workspace(self, n, d, elem_size, device):
        return self.capi.alloc_workspace(n, d, elem_size, device)

    def plan_unique(self, ids, height, world, ws, want_elem_bounds):
        return self.capi.plan_unique_rows(ids, height, world, ws, want_elem_bounds=want_elem_bounds)

    def expand_unique(self, unique_values, n, plan_h, ws):
        return self.capi.expand_unique(unique_values, n, plan_h, ws)

    def gather_sorted(self, value, plan_h, ws):
        return self.capi.gather_sorted(value, plan_h, ws)

    def merge_by_plan(self, ids, value, plan_h, ws):
        _, merged, _ = self.capi.merge_add(ids, value, plan_h, workspace=ws, flags=self.capi.PLAN_READY)
        return merged

    def lookup(self, shard, local_rows):
        return self.capi.lookup_table(shard, local_rows)

    def sgd(self, shard, rows, grads, lr, deterministic, ws=None, plan_ready=False):
        flags = (self.capi.DETERMINISTIC if deterministic else 0) | (self.capi.PLAN_READY if plan_ready else 0)
        self.capi.sparse_sgd(shard, rows, grads, lr, workspace=ws, flags=flags)

    def adam(self, shard, m1, m2, rows, grads, lr, b1p, b2p, beta1, beta2, eps, deterministic, ws=None,
             plan_ready=False):
        flags = (self.capi.DETERMINISTIC if deterministic else 0) | (self.capi.PLAN_READY if plan_ready else 0)
        self.capi.sparse_adam(shard, m1, m2, rows, grads, lr, b1p, b2p, beta1, beta2, eps, workspace=ws,
                              flags=flags)


def all_to_all_rows(out, inp, out_splits, in_splits, group=None):
    """Variable-size all-to-all along dim 0.  NCCL: one grouped send/recv; gloo: isend/irecv pairs."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        out.copy_(inp)
        return
    if dist.get_backend(group) == "nccl":
        dist.all_to_all_single(out, inp, output_split_sizes=list(out_splits),
                               input_split_sizes=list(in_splits), group=group)
        return
    rank = dist.get_rank(group)
    outs = list(out.split(list(out_splits), dim=0))
    ins = list(inp.split(list(in_splits), dim=0))
    outs[rank].copy_(ins[rank])
    reqs = []
    for peer in range(world):
        if peer == rank:
            continue
        if ins[peer].numel():
            reqs.append(dist.isend(ins[peer].contiguous(), peer, group=group))
        if outs[peer].numel():
            reqs.append(dist.irecv(outs[peer], peer, group=group))
    for r in reqs:
        r.wait()


class _Phase:
    """with table._phase("name"): ... records a CUDA-event pair when table.phase_events is a list."""

    def __init__(self, table, name):
        self.t, self.name = table, name

    def __enter__(self):
        if self.t.phase_events is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if self.t.phase_events is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            self.t.phase_events.append((self.name, self.e0, e1))
        return False


class ShardedTable:
    """This rank's shard of a [height, d] table plus the exchange logic of one training step."""

    def __init__(self, height, d, ops, device, dtype=torch.float32, group=None, init=None, optimizer="sgd"):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.height, self.d = int(height), int(d)
        self.hs = shard_height(height, self.world)
        self.plan_h = self.world * self.hs if self.world > 1 else self.height
        self.ops, self.device, self.dtype = ops, device, dtype
        self.shard = torch.zeros((self.hs, d), dtype=dtype, device=device)
        if init is not None:          # init(global_row_ids[int64]) -> [len, d] values, for parity tests
            rows = torch.arange(self.rank, self.height, self.world, dtype=torch.int64)
            self.shard[: rows.numel()] = init(rows).to(device=device, dtype=dtype)
        self.optimizer = optimizer
        if optimizer == "adam":
            self.m1 = torch.zeros_like(self.shard)
            self.m2 = torch.zeros_like(self.shard)
        self.bytes_sent = 0
        self.peer = False
        self.phase_events = None      # set to [] to record (name, start_event, end_event) per phase

    def _phase(self, name):
        return _Phase(self, name)

    def enable_engine(self, max_ids, max_recv):
        """Set-up (once, fast mode, float32): create this rank's step engine (b200emb_shard_create), exchange
        the CUDA-IPC handles of the ranks' blocks and connect.  max_ids bounds the ids of one step on this rank,
        max_recv the unique rows all ranks together request from this owner in one step; exceeding either raises
        at the next check_status().  The shard (and Adam moments) stay the tensors of this object."""
        capi = self.ops.capi
        if self.dtype != torch.float32:
            return False
        self.engine = capi.ShardEngine(self.world, self.rank, self.height, self.d, self.shard, int(max_ids),
                                       int(max_recv), self.optimizer, getattr(self, "m1", None),
                                       getattr(self, "m2", None))
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, self.engine.handle, group=self.group)
            self.engine.connect(handles)
        else:
            capi.ShardEngine.connect_local([self.engine])
        self._slot = -1
        self.peer = True
        return True

    def check_status(self):
        """Raise if a kernel of the engine path flagged a rank time-out or a capacity overflow.  Reads one device
        word (a synchronisation): call it where the training loop synchronises anyway."""
        if self.peer:
            return self.engine.check_status()
        return 0

    # ------------------------------------------------------------------ forward
    def prepare(self, ids, ws, deterministic=False, stream=None, after=None):
        """The ids-only part of lookup(): sort plan + de-duplication in `ws`, the split sizes of every rank
        (one all_gather) and their copy to the host.  Needs nothing but the ids, so a training loop can run
        it for batch i+1 on a side `stream` while batch i is still in flight (a data-loader prefetch); the
        host then finds the sizes already there instead of stalling the device queue on a mid-step read."""
        if self.peer and ids.is_cuda and not deterministic:
            self._slot = (self._slot + 1) % 4
            if stream is not None:
                stream.wait_stream(after if after is not None else torch.cuda.current_stream(ids.device))
            self.engine.prepare(self._slot, ids, stream=stream)
            event = torch.cuda.Event()
            event.record(stream if stream is not None else torch.cuda.current_stream(ids.device))
            return {"ids": ids, "ws": ws, "deterministic": False, "slot": self._slot, "event": event,
                    "stream": stream, "engine": True}
        cur = torch.cuda.current_stream(ids.device) if ids.is_cuda else None
        use = stream if (stream is not None and ids.is_cuda) else None
        if use is not None:
            use.wait_stream(cur)             # ws / ids of this batch are free of earlier steps' work
        with (torch.cuda.stream(use) if use is not None else contextlib.nullcontext()):
            with self._phase("plan+dedup"):
                ukeys, n_unique, bounds, ebounds, local = self.ops.plan_unique(ids, self.height, self.world, ws,
                                                                               deterministic)
            mine = bounds if ebounds is None else torch.cat([bounds, ebounds])
            with self._phase("split sizes (all_gather)"):
                if self.world > 1:
                    flat = torch.empty(self.world * mine.numel(), dtype=mine.dtype, device=mine.device)
                    dist.all_gather_into_tensor(flat, mine, group=self.group)
                    table_d = flat.view(self.world, mine.numel())
                else:
                    table_d = mine.reshape(1, -1)
            event = None
            if ids.is_cuda:
                slot = self._host_slot = (getattr(self, "_host_slot", -1) + 1) % 4
                if not hasattr(self, "_host_tables"):
                    self._host_tables = [None] * 4
                if self._host_tables[slot] is None or self._host_tables[slot].shape != table_d.shape:
                    self._host_tables[slot] = torch.empty(table_d.shape, dtype=table_d.dtype).pin_memory()
                table = self._host_tables[slot]
                table.copy_(table_d, non_blocking=True)
                event = torch.cuda.Event()
                event.record()
            else:
                table = table_d
        if use is not None:
            for t in (ukeys, n_unique, bounds, ebounds, local, table_d):
                if t is not None:
                    t.record_stream(cur)
        return {"ids": ids, "ws": ws, "deterministic": deterministic, "bounds": bounds, "ebounds": ebounds,
                "local": local, "table_d": table_d, "table": table, "event": event, "stream": use}

    def lookup(self, ids, ws, deterministic=False, prep=None, owner_ws=None, owner_stream=None):
        """ids: int64 [n] on this rank.  Returns (out [n, d], ctx for update()).  `prep` = the result of
        prepare() on the same (ids, ws) if it was issued ahead of time.  (owner_ws / owner_stream: kept for
        callers of the earlier interface; the engine owns its owner-side workspace and stream.)"""
        n = ids.numel()
        if prep is None:
            prep = self.prepare(ids, ws, deterministic)
        assert prep["ids"] is ids and prep["ws"] is ws and prep["deterministic"] == deterministic
        if prep.get("engine"):
            if prep["stream"] is not None:
                torch.cuda.current_stream(ids.device).wait_event(prep["event"])
            with self._phase("engine lookup (send ids, collect, owner gather, expand)"):
                out = self.engine.lookup(prep["slot"])
            ctx = {"engine": True, "slot": prep["slot"], "ids": ids, "n": n, "fused": True}
            return out, ctx
        bounds, ebounds, local, table_d = prep["bounds"], prep["ebounds"], prep["local"], prep["table_d"]
        with self._phase("split sizes (host read)"):
            if prep["event"] is not None:
                prep["event"].synchronize()
                if prep["stream"] is not None:
                    torch.cuda.current_stream(ids.device).wait_event(prep["event"])
            table = prep["table"].numpy().copy()      # G x (G+1) integers (numpy: cheaper than tensor ops here)
        g1 = self.world + 1
        ub = table[:, :g1]
        send = (ub[self.rank, 1:] - ub[self.rank, :-1]).tolist()
        recv = (ub[:, self.rank + 1] - ub[:, self.rank]).tolist()
        u = int(ub[self.rank, self.world])
        recv_local = torch.empty(sum(recv), dtype=torch.int64, device=ids.device)
        with self._phase("id all-to-all"):
            all_to_all_rows(recv_local, local[:u], recv, send, self.group)
        with self._phase("owner gather"):
            rows = self.ops.lookup(self.shard, recv_local)
        uvals = torch.empty((u, self.d), dtype=self.dtype, device=ids.device)
        with self._phase("row all-to-all"):
            all_to_all_rows(uvals, rows, send, recv, self.group)
        with self._phase("expand to batch order"):
            out = self.ops.expand_unique(uvals, n, self.plan_h, ws)
        self.bytes_sent += (u - send[self.rank]) * 8 + (sum(recv) - recv[self.rank]) * self.d * rows.element_size()
        ctx = {"ids": ids, "n": n, "u": u, "send": send, "recv": recv, "recv_local": recv_local, "ws": ws,
               "fused": False, "bounds": bounds, "ub": ub}
        if ebounds is not None:
            eb = table[:, g1:]
            ctx["esend"] = (eb[self.rank, 1:] - eb[self.rank, :-1]).tolist()
            ctx["erecv"] = (eb[:, self.rank + 1] - eb[:, self.rank]).tolist()
        return out, ctx

    # ------------------------------------------------------------------ backward + update
    def update(self, ctx, d_out, lr, deterministic=False, adam_state=None, owner_ws=None):
        if ctx.get("engine"):
            with self._phase("engine update (local MergeAdd -> owners, fused owner update)"):
                if self.optimizer == "adam":
                    b1p, b2p, beta1, beta2, eps = adam_state
                    self.engine.update(ctx["slot"], d_out, lr, b1p, b2p, beta1, beta2, eps)
                else:
                    self.engine.update(ctx["slot"], d_out, lr)
            return
        ids, ws = ctx["ids"], ctx["ws"]
        if not deterministic:
            with self._phase("local MergeAdd"):
                merged = self.ops.merge_by_plan(ids, d_out, self.plan_h, ws)[: ctx["u"]]
            grads = torch.empty((sum(ctx["recv"]), self.d), dtype=self.dtype, device=d_out.device)
            with self._phase("grad all-to-all"):
                all_to_all_rows(grads, merged, ctx["recv"], ctx["send"], self.group)
            rows = ctx["recv_local"]
            self.bytes_sent += (ctx["u"] - ctx["send"][self.rank]) * self.d * merged.element_size()
        else:
            # ship every occurrence: row ids and gradient rows in plan order, element-level splits
            sorted_grads = self.ops.gather_sorted(d_out, self.plan_h, ws)
            ids_f64 = ids.view(torch.float64).reshape(-1, 1)          # bit copy of the int64 ids
            sorted_ids = self.ops.gather_sorted(ids_f64, self.plan_h, ws).view(torch.int64).reshape(-1)
            esend, erecv = ctx["esend"], ctx["erecv"]
            nvalid = sum(esend)
            rows = torch.empty(sum(erecv), dtype=torch.int64, device=d_out.device)
            all_to_all_rows(rows, torch.div(sorted_ids[:nvalid], self.world, rounding_mode="floor")
                            if self.world > 1 else sorted_ids[:nvalid], erecv, esend, self.group)
            grads = torch.empty((sum(erecv), self.d), dtype=self.dtype, device=d_out.device)
            all_to_all_rows(grads, sorted_grads[:nvalid], erecv, esend, self.group)
        if rows.numel() == 0:
            return
        with self._phase("owner fused update"):
            if self.optimizer == "adam":
                b1p, b2p, beta1, beta2, eps = adam_state
                self.ops.adam(self.shard, self.m1, self.m2, rows, grads, lr, b1p, b2p, beta1, beta2, eps,
                              deterministic, owner_ws)
            else:
                self.ops.sgd(self.shard, rows, grads, lr, deterministic, owner_ws)

    def gather_full_table(self):
        """Debug/parity helper: the full [height, d] table on every rank (interleaved shards)."""
        if self.world == 1:
            return self.shard[: self.height].clone()
        parts = [torch.empty_like(self.shard) for _ in range(self.world)]
        dist.all_gather(parts, self.shard, group=self.group)
        full = torch.stack(parts, dim=1).reshape(self.world * self.hs, self.d)
        return full[: self.height]
