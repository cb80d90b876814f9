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
The only host synchronisation per step is the read of the split sizes (G x (G+1) integers).

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

    # fused compute + exchange: the kernels store straight into peer GPUs' buffers over NVLink
    def lookup_scatter(self, shard, local_rows, seg_start, seg_dst):
        self.capi.lookup_table_scatter(shard, local_rows, seg_start, seg_dst)

    def ids_scatter(self, local_rows, n_upper, bounds, dst):
        self.capi.segment_scatter_i64(local_rows, n_upper, bounds, dst)

    def merge_scatter(self, ids, value, plan_h, shard_bounds, shard_dst, ws):
        self.capi.merge_add_scatter(ids, value, plan_h, shard_bounds, shard_dst, ws,
                                    flags=self.capi.PLAN_READY | self.capi.UNIQUE_READY)

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

    def enable_peer_exchange(self, cap_unique, cap_recv):
        """Set-up (once): exchange CUDA-IPC handles of two receive buffers per rank so that the owner's
        gather kernel and the requester's MergeAdd kernel can store straight into the other GPUs' memory
        (fused compute + all-to-all over NVLink).  Steps whose sizes exceed the capacities fall back to NCCL."""
        if self.world == 1:
            return False
        capi = self.ops.capi
        item = torch.empty(0, dtype=self.dtype).element_size()
        self.cap_u, self.cap_r = int(cap_unique), int(cap_recv)
        self.peer_u = capi.PeerBuffer(self.cap_u * self.d * item)
        self.peer_g = capi.PeerBuffer(self.cap_r * self.d * item)
        self.peer_i = capi.PeerBuffer(self.cap_r * 8)
        self.peer_f = capi.PeerBuffer(max(self.world * 8, 256))        # barrier flags: uint64[world]
        self.peer_f.tensor((max(self.world, 32),), torch.int64).zero_()
        torch.cuda.synchronize(self.device)
        handles = [None] * self.world
        dist.all_gather_object(handles, (self.peer_u.handle, self.peer_g.handle, self.peer_i.handle,
                                         self.peer_f.handle), group=self.group)
        self.peer_u_ptr, self.peer_g_ptr, self.peer_i_ptr, peer_f_ptr = [], [], [], []
        for r, (hu, hg, hi, hf) in enumerate(handles):
            if r == self.rank:
                self.peer_u_ptr.append(self.peer_u.ptr)
                self.peer_g_ptr.append(self.peer_g.ptr)
                self.peer_i_ptr.append(self.peer_i.ptr)
                peer_f_ptr.append(self.peer_f.ptr)
            else:
                self.peer_u_ptr.append(capi.peer_open(hu))
                self.peer_g_ptr.append(capi.peer_open(hg))
                self.peer_i_ptr.append(capi.peer_open(hi))
                peer_f_ptr.append(capi.peer_open(hf))
        self.peer_f_ptr_t = torch.tensor(peer_f_ptr, dtype=torch.int64, device=self.device)
        # split-size all-gather over peer memory: two tables (even / odd step) of world rows x 2(world+1)
        # int64, and a flag array of its own
        self.ag_cols = 2 * (self.world + 1)
        tbytes = self.world * self.ag_cols * 8
        self.peer_t = [capi.PeerBuffer(max(tbytes, 256)) for _ in range(2)]
        self.peer_af = capi.PeerBuffer(max(self.world * 8, 256))
        self.peer_af.tensor((max(self.world, 32),), torch.int64).zero_()
        torch.cuda.synchronize(self.device)
        h2 = [None] * self.world
        dist.all_gather_object(h2, (self.peer_t[0].handle, self.peer_t[1].handle, self.peer_af.handle),
                               group=self.group)
        t_ptr, af_ptr = [[], []], []
        for r, (h0, h1, ha) in enumerate(h2):
            mine = r == self.rank
            t_ptr[0].append(self.peer_t[0].ptr if mine else capi.peer_open(h0))
            t_ptr[1].append(self.peer_t[1].ptr if mine else capi.peer_open(h1))
            af_ptr.append(self.peer_af.ptr if mine else capi.peer_open(ha))
        self.peer_t_ptr_t = [torch.tensor(p, dtype=torch.int64, device=self.device) for p in t_ptr]
        self.peer_af_ptr_t = torch.tensor(af_ptr, dtype=torch.int64, device=self.device)
        self.ag_epoch = 0
        self._slots = [None] * 4          # preallocated outputs of prepare(), used round-robin
        self._slot = -1
        self.barrier_epoch = 0
        self.barrier_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        # collective fallback (A/B switch): a 4-byte all_reduce
        self.nccl_barrier = os.environ.get("B200EMB_NCCL_BARRIER") is not None
        self.peer_i_ptr_t = torch.tensor(self.peer_i_ptr, dtype=torch.int64, device=self.device)
        self.peer_u_ptr_t = torch.tensor(self.peer_u_ptr, dtype=torch.int64, device=self.device)
        self.peer_g_ptr_t = torch.tensor(self.peer_g_ptr, dtype=torch.int64, device=self.device)
        self.sync_word = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.row_bytes = self.d * item
        self.peer = True
        return True

    def _device_barrier(self):
        # stream-ordered cross-GPU barrier: every rank's preceding kernels (peer stores) have completed
        if self.nccl_barrier:
            dist.all_reduce(self.sync_word, group=self.group)
            return
        self.barrier_epoch += 1
        self.ops.capi.peer_barrier(self.peer_f_ptr_t, self.world, self.rank, self.barrier_epoch,
                                   self.barrier_status)

    # ------------------------------------------------------------------ forward
    def prepare(self, ids, ws, deterministic=False, stream=None, after=None):
        """The ids-only part of lookup(): sort plan + de-duplication in `ws`, the split sizes of every rank
        (one all_gather) and their copy to the host.  Needs nothing but the ids, so a training loop can run
        it for batch i+1 on a side `stream` while batch i is still in flight (a data-loader prefetch); the
        host then finds the sizes already there instead of stalling the device queue on a mid-step read."""
        if self.peer and ids.is_cuda and not self.nccl_barrier:
            return self._prepare_peer(ids, ws, deterministic, stream, after)
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
            maps = None
            if self.peer and ids.is_cuda:
                # scatter maps of the fused exchange, computed on the device from the gathered split table
                m = self.ops.capi.exchange_maps(table_d, self.rank, self.peer_i_ptr_t, self.peer_u_ptr_t,
                                                self.peer_g_ptr_t, self.row_bytes)
                w = self.world
                maps = {"all": m, "id_dst": m[0, :w], "seg_start": m[1], "seg_dst": m[2, :w], "shard_dst": m[3, :w]}
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
            for t in (ukeys, n_unique, bounds, ebounds, local, table_d, maps["all"] if maps else None):
                if t is not None:
                    t.record_stream(cur)
        return {"ids": ids, "ws": ws, "deterministic": deterministic, "bounds": bounds, "ebounds": ebounds,
                "local": local, "table_d": table_d, "table": table, "event": event, "stream": use, "maps": maps}

    def _prepare_peer(self, ids, ws, deterministic, stream, after):
        """prepare() when the peer buffers exist: three library calls on `stream` (plan + de-duplication,
        all-gather of the split sizes over peer memory incl. the copy to pinned host memory, scatter maps),
        all outputs preallocated — the host spends a few tens of microseconds here."""
        capi, w, n = self.ops.capi, self.world, ids.numel()
        self._slot = (self._slot + 1) % len(self._slots)
        sl = self._slots[self._slot]
        if sl is None or sl["ukeys"].numel() < max(n, 1):
            dev = ids.device
            sl = self._slots[self._slot] = {
                "ukeys": torch.empty(max(n, 1), dtype=torch.int64, device=dev),
                "local": torch.empty(max(n, 1), dtype=torch.int64, device=dev),
                "n_unique": torch.zeros(1, dtype=torch.int64, device=dev),
                "row": torch.zeros(self.ag_cols, dtype=torch.int64, device=dev),     # bounds | elem bounds
                "maps": torch.empty((4, w + 1), dtype=torch.int64, device=dev),
                "host": torch.empty((w, self.ag_cols), dtype=torch.int64).pin_memory(),
                "event": torch.cuda.Event(),
            }
        if stream is not None:
            stream.wait_stream(after if after is not None else torch.cuda.current_stream(ids.device))
        bounds, ebounds = sl["row"][: w + 1], sl["row"][w + 1:]
        capi.plan_unique_rows(ids, self.height, w, ws, want_elem_bounds=deterministic,
                              out=(sl["ukeys"], sl["n_unique"], bounds, ebounds, sl["local"]), stream=stream)
        self.ag_epoch += 1
        par = self.ag_epoch & 1
        local_table = self.peer_t[par].tensor((w, self.ag_cols), torch.int64)
        capi.peer_allgather_i64(sl["row"], self.peer_t_ptr_t[par], local_table, self.peer_af_ptr_t, w, self.rank,
                                self.ag_epoch, host_copy=sl["host"], status=self.barrier_status, stream=stream)
        m = capi.exchange_maps(local_table, self.rank, self.peer_i_ptr_t, self.peer_u_ptr_t, self.peer_g_ptr_t,
                               self.row_bytes, out=sl["maps"], stream=stream)
        sl["event"].record(stream if stream is not None else torch.cuda.current_stream(ids.device))
        maps = {"all": m, "id_dst": m[0, :w], "seg_start": m[1], "seg_dst": m[2, :w], "shard_dst": m[3, :w]}
        g1 = w + 1
        return {"ids": ids, "ws": ws, "deterministic": deterministic, "bounds": bounds,
                "ebounds": ebounds if deterministic else None, "local": sl["local"], "table_d": local_table,
                "table": sl["host"], "event": sl["event"], "stream": stream, "maps": maps,
                "cols": g1 * (2 if deterministic else 1)}

    def lookup(self, ids, ws, deterministic=False, prep=None, owner_ws=None, owner_stream=None):
        """ids: int64 [n] on this rank.  Returns (out [n, d], ctx for update()).  `prep` = the result of
        prepare() on the same (ids, ws) if it was issued ahead of time.  With `owner_ws` (the workspace
        update() will get) the owner-side sort plan of the received row ids is built right after the id
        exchange — on `owner_stream` if given — instead of on update()'s critical path."""
        n = ids.numel()
        if prep is None:
            prep = self.prepare(ids, ws, deterministic)
        assert prep["ids"] is ids and prep["ws"] is ws and prep["deterministic"] == deterministic
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
        fused = (self.peer and not deterministic and int(ub[:, self.world].max()) <= self.cap_u and
                 int((ub[:, 1:] - ub[:, :-1]).sum(axis=0).max()) <= self.cap_r)
        if fused:
            maps = prep["maps"]
            with self._phase("id scatter to owners (peer stores)"):
                self.ops.ids_scatter(local, n, bounds, maps["id_dst"])
            with self._phase("barrier"):
                self._device_barrier()
            # private copy: a faster peer may start the NEXT step's id scatter into peer_i while this rank's
            # owner update still needs these ids (nothing between the two orders them)
            recv_local = self.peer_i.tensor((sum(recv),), torch.int64).clone()
        else:
            recv_local = torch.empty(sum(recv), dtype=torch.int64, device=ids.device)
            with self._phase("id all-to-all"):
                all_to_all_rows(recv_local, local[:u], recv, send, self.group)
        # the owner-side plan of the rows this rank will update needs their ids only: start it now
        owner_plan = None
        if (owner_ws is not None and not deterministic and recv_local.numel() > 0 and recv_local.is_cuda and
                owner_ws.numel() >= self.ops.capi.workspace_bytes(recv_local.numel(), self.d,
                                                                  self.shard.element_size())):
            cur = torch.cuda.current_stream(ids.device)
            if owner_stream is not None:
                owner_stream.wait_stream(cur)
            if self.phase_events is not None:
                with (torch.cuda.stream(owner_stream) if owner_stream is not None else contextlib.nullcontext()):
                    with self._phase("owner sort plan (ids only)"):
                        self.ops.capi.plan_rows(recv_local, self.hs, owner_ws)
            else:
                self.ops.capi.plan_rows(recv_local, self.hs, owner_ws, stream=owner_stream)
            owner_plan = self._owner_events[self._owner_slot] if hasattr(self, "_owner_events") else None
            if owner_plan is None:
                self._owner_events = [torch.cuda.Event() for _ in range(4)]
                self._owner_slot = 0
                owner_plan = self._owner_events[0]
            self._owner_slot = (self._owner_slot + 1) % 4
            owner_plan.record(owner_stream if owner_stream is not None else cur)
            if owner_stream is not None:
                recv_local.record_stream(owner_stream)
        if fused:
            # owner side: gather from the shard straight into every requester's unique-row buffer
            with self._phase("owner gather + peer stores (fused row exchange)"):
                self.ops.lookup_scatter(self.shard, recv_local, maps["seg_start"], maps["seg_dst"])
            with self._phase("barrier"):
                self._device_barrier()
            uvals = self.peer_u.tensor((u, self.d), self.dtype)
            rows = uvals
        else:
            with self._phase("owner gather"):
                rows = self.ops.lookup(self.shard, recv_local)
            uvals = torch.empty((u, self.d), dtype=self.dtype, device=ids.device)
            with self._phase("row all-to-all"):
                all_to_all_rows(uvals, rows, send, recv, self.group)
        with self._phase("expand to batch order"):
            out = self.ops.expand_unique(uvals, n, self.plan_h, ws)
        self.bytes_sent += (u - send[self.rank]) * 8 + (sum(recv) - recv[self.rank]) * self.d * rows.element_size()
        ctx = {"ids": ids, "n": n, "u": u, "send": send, "recv": recv, "recv_local": recv_local, "ws": ws,
               "fused": fused, "bounds": bounds, "ub": ub, "maps": prep["maps"] if fused else None,
               "owner_plan": owner_plan, "owner_ws": owner_ws}
        if ebounds is not None:
            eb = table[:, g1:]
            ctx["esend"] = (eb[self.rank, 1:] - eb[self.rank, :-1]).tolist()
            ctx["erecv"] = (eb[:, self.rank + 1] - eb[:, self.rank]).tolist()
        return out, ctx

    # ------------------------------------------------------------------ backward + update
    def update(self, ctx, d_out, lr, deterministic=False, adam_state=None, owner_ws=None):
        ids, ws = ctx["ids"], ctx["ws"]
        if not deterministic and ctx.get("fused"):
            # requester side: the local duplicate reduction stores every merged row straight into its owner's
            # receive buffer, at this rank's slot (ranks before it come first: reference gather-concat order)
            with self._phase("local MergeAdd + peer stores (fused grad exchange)"):
                self.ops.merge_scatter(ids, d_out, self.plan_h, ctx["bounds"], ctx["maps"]["shard_dst"], ws)
            with self._phase("barrier"):
                self._device_barrier()
            rows = ctx["recv_local"]
            grads = self.peer_g.tensor((sum(ctx["recv"]), self.d), self.dtype)
            self.bytes_sent += (ctx["u"] - ctx["send"][self.rank]) * self.row_bytes
        elif not deterministic:
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
        extra = {}
        if ctx.get("owner_plan") is not None and owner_ws is ctx.get("owner_ws") and rows is ctx["recv_local"]:
            torch.cuda.current_stream(d_out.device).wait_event(ctx["owner_plan"])
            extra = {"plan_ready": True}          # lookup() already sorted these row ids into owner_ws
        with self._phase("owner fused update"):
            if self.optimizer == "adam":
                b1p, b2p, beta1, beta2, eps = adam_state
                self.ops.adam(self.shard, self.m1, self.m2, rows, grads, lr, b1p, b2p, beta1, beta2, eps,
                              deterministic, owner_ws, **extra)
            else:
                self.ops.sgd(self.shard, rows, grads, lr, deterministic, owner_ws, **extra)

    def gather_full_table(self):
        """Debug/parity helper: the full [height, d] table on every rank (interleaved shards)."""
        if self.world == 1:
            return self.shard[: self.height].clone()
        parts = [torch.empty_like(self.shard) for _ in range(self.world)]
        dist.all_gather(parts, self.shard, group=self.group)
        full = torch.stack(parts, dim=1).reshape(self.world * self.hs, self.d)
        return full[: self.height]
