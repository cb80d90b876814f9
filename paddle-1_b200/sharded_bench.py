"""bench.py --gpus N (N > 1): BASELINE.json configs[3] — 100M x 64 fp32 table row-sharded over the N
GPUs of one box (owner = id % N), Zipf(1.1) ids, id / row / gradient all-to-alls over NVLink, sparse SGD.
Launched by torchrun, one rank per GPU.  Weak scaling: every GPU brings its own batch."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

from . import capi, sharded, synth

METRIC = "embedding lookups+updates/sec"
UNIT = "ids/s"
N_ROTATE = 4
NVLINK_GBS = 770.0   # measured peer-copy bandwidth per direction per GPU (B200_PROFILING.md)


def run(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != args.gpus:
        raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    capi.lib()
    v, d, n_base, a, opt = synth.CONFIGS["C4_sharded_sgd"]
    n = int(getattr(args, "ids_per_gpu", 0) or n_base * 16)
    ops = sharded.CudaOps(capi)
    tab = sharded.ShardedTable(v, d, ops, dev)
    fused = os.environ.get("B200EMB_NO_PEER") is None and tab.enable_peer_exchange(n, 2 * n)
    # random-init shard, in slabs to keep the peak low
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    for lo in range(0, tab.hs, 1 << 22):
        hi = min(tab.hs, lo + (1 << 22))
        tab.shard[lo:hi].uniform_(-0.05, 0.05, generator=g)
    lr = torch.tensor([0.01], dtype=torch.float32, device=dev)
    z = synth.ZipfIds(v, a, seed=rank)
    ids_h = [torch.from_numpy(z.sample(n)).pin_memory() for _ in range(N_ROTATE)]
    del z
    dout_h = [torch.from_numpy(synth.grad_values(n, d, seed=4321 + rank * 16 + b)).pin_memory() for b in range(N_ROTATE)]
    ids = [t.to(dev) for t in ids_h]
    dout = [t.to(dev) for t in dout_h]
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(N_ROTATE)]
    owner_ws = capi.alloc_workspace_upto(2 * n, d, 4, dev)
    owner_cap = 2 * n
    state = {"owner_ws": owner_ws, "cap": owner_cap}

    # The ids-only part of a step (sort plan + de-duplication + split sizes) of batch i+1 is issued on a
    # side stream BEFORE the main-stream work of batch i is queued — the prefetch a data loader gives for
    # free — so the host finds the sizes of the next batch already copied and never stalls the device queue.
    side = torch.cuda.Stream(dev)
    side2 = torch.cuda.Stream(dev)
    main_s = torch.cuda.current_stream(dev)
    prefetch = os.environ.get("B200EMB_NO_PREFETCH") is None

    def step(b, prep=None, next_b=None, ids_b=None, dout_b=None):
        i = ids[b] if ids_b is None else ids_b
        g_ = dout[b] if dout_b is None else dout_b
        nxt = tab.prepare(ids[next_b], ws[next_b], stream=side, after=main_s) if (prefetch and next_b is not None) else None
        out, ctx = tab.lookup(i, ws[b], prep=prep, owner_ws=state["owner_ws"] if prefetch else None,
                              owner_stream=side2)
        need = sum(ctx["recv"])
        if need > state["cap"]:
            state["cap"] = 2 * need
            state["owner_ws"] = capi.alloc_workspace_upto(state["cap"], d, 4, dev)
        tab.update(ctx, g_, lr, owner_ws=state["owner_ws"])
        return out, ctx, nxt

    def run_steps(k):
        """k consecutive steps; exactly k plans are issued inside the call (the first one inline-ahead,
        none after the last step)."""
        prep = tab.prepare(ids[0], ws[0], stream=side, after=main_s) if prefetch else None
        out = ctx = None
        for i in range(k):
            nb = (i + 1) % N_ROTATE if i + 1 < k else None
            out, ctx, prep = step(i % N_ROTATE, prep, nb)
        return out, ctx

    # at least 10 untimed steps: the first ones map peer pages, fill the allocator's pools and ramp the clocks
    warm = max(args.warmup, 10)
    out, ctx = run_steps(warm)
    torch.cuda.synchronize()
    u_local, recv_rows = ctx["u"], sum(ctx["recv"])
    tab.bytes_sent = 0
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if os.environ.get("B200EMB_PROFILE_HOST") and rank == 0:      # developer switch: where the host time goes
        import cProfile, pstats, sys
        pr = cProfile.Profile()
        pr.enable()
        run_steps(args.steps)
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats(os.environ.get("B200EMB_PROFILE_SORT", "tottime")).print_stats(45)
    else:
        run_steps(args.steps)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    sent = torch.tensor([float(tab.bytes_sent)], device=dev)
    dist.all_reduce(sent, op=dist.ReduceOp.MAX)
    dist.barrier()

    # per-phase device times (rank 0), a few extra untimed steps
    phases = {}
    tab.phase_events = []
    run_steps(4)
    torch.cuda.synchronize()
    for name, a0, a1 in tab.phase_events:
        phases[name] = phases.get(name, 0.0) + a0.elapsed_time(a1) / 4 * 1e3
    tab.phase_events = None
    dist.barrier()

    # end to end: host batches in, rows out, every step
    e2e_steps = max(5, min(args.steps, 30))
    ids_d = torch.empty(n, dtype=torch.int64, device=dev)
    dout_d = torch.empty((n, d), dtype=torch.float32, device=dev)
    out_h = torch.empty((n, d), dtype=torch.float32).pin_memory()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        b = i % N_ROTATE
        ids_d.copy_(ids_h[b], non_blocking=True)
        dout_d.copy_(dout_h[b], non_blocking=True)
        out, _, _ = step(b, ids_b=ids_d, dout_b=dout_d)
        out_h.copy_(out, non_blocking=True)
    torch.cuda.synchronize()
    el = torch.tensor([time.perf_counter() - t0], device=dev)
    dist.all_reduce(el, op=dist.ReduceOp.MAX)
    el = float(el.item())

    if rank == 0:
        try:
            with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
                peak = float(json.load(f)["hbm_gbs"])
        except Exception:
            peak = 6650.0
        ms_per_step = ms / args.steps
        value = world * n * args.steps / (ms * 1e-3)
        # per-GPU algorithmic HBM bytes: requester side (ids, expand write, dOut read) + owner side
        # (gather read+write of received rows, fused update of received rows)
        b_req = n * (8 + 4 * d) + n * (8 + 4 * d)
        b_own = recv_rows * (8 + 2 * 4 * d) + recv_rows * (8 + 4 * d) + recv_rows * 2 * 4 * d
        hbm = (b_req + b_own) / (ms_per_step * 1e-3) / 1e9
        nv = float(sent.item()) / args.steps / (ms_per_step * 1e-3) / 1e9
        # this library's kernels per step on the fused path — prepare: sort plan, head scan (2), inverse map,
        # shard split, split-size all-gather, scatter maps; lookup: id scatter, barrier, owner-side sort plan,
        # owner gather, barrier, expand; update: MergeAdd + scatter, barrier, owner update
        launches = (capi.plan_kernel_count(n, tab.plan_h) + capi.HEAD_SCAN_KERNELS + 4) + 2 + \
                   capi.plan_kernel_count(recv_rows, tab.hs) + 3 + 3
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[3]: synthetic 100Mx64 fp32 table row-sharded over %d B200 (owner = id %% G), "
                                   "Zipf(1.1) ids, id/row/grad all-to-all over NVLink (NCCL), sparse SGD" % world,
                       "table": [v, d], "ids_per_gpu_per_step": n, "unique_ids_per_gpu": u_local,
                       "rows_received_per_gpu": recv_rows, "optimizer": opt, "mode": "fast, dedup before send, " + ("gather and MergeAdd kernels store into peer GPUs over NVLink (no NCCL all-to-all for rows/grads)" if fused else "NCCL all-to-all for rows/grads"),
                       "l2": "inputs rotate over %d batch buffers per GPU (%.0f MB) > 126 MB L2" % (N_ROTATE, N_ROTATE * n * (8 + 8 * d) / 1e6),
                       "launch": ("the ids-only part of step i+1 (sort plan, de-duplication, split-size all_gather) runs on a "
                                  "side stream during step i (data-loader prefetch); every step still does one of each"
                                  if prefetch else "no prefetch: every step reads its split sizes on the host mid-step")},
            "roofline": {"bound": "hbm", "achieved": hbm, "peak": peak, "unit": "GB/s", "frac": hbm / peak,
                         "traffic": None, "kernel": "whole sharded step, per GPU (requester + owner side)",
                         "nvlink": {"achieved_GBs_per_direction": nv, "peak": NVLINK_GBS, "frac": nv / NVLINK_GBS}},
            "e2e": {"value": world * n * e2e_steps / el, "unit": UNIT, "h2d_bytes_per_step": world * (n * 8 + n * d * 4),
                    "d2h_bytes_per_step": world * n * d * 4, "ms_per_step": el / e2e_steps * 1e3, "steps": e2e_steps},
            "phases_us": {k: round(x, 1) for k, x in phases.items()},
            "gpu_launches": launches * args.steps,
        }
        print(json.dumps(line))
    dist.barrier()
    dist.destroy_process_group()
