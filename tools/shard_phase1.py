#!/usr/bin/env python
"""Developer tool: per-phase times of the sharded engine with world = 1 on one GPU (every "peer" store is local):
what the step costs without NVLink, to set beside the multi-GPU phase times of bench.py --gpus N."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_sharded as bs  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
torch.cuda.set_device(0)
dev = torch.device("cuda:0")
dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
pkg = load_package()
capi, synth = pkg.capi, pkg.synth
v, d = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000, 64
shard = torch.empty((v, d), device=dev).uniform_(-0.05, 0.05)
zipf = synth.ZipfIds(v, 1.1, seed=0)
lr = torch.tensor([0.01], device=dev)
for n in [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["106496", "1703936"])]:
    wl = bs.Workload(pkg, zipf, n, d, 0, dev, pinned=False)
    cap, _ = bs.recv_capacity(wl.ids_np, 1, 0)
    eng = capi.ShardEngine(1, 0, v, d, shard, n, int(cap * 1.25) + 1024)
    capi.ShardEngine.connect_local([eng])
    ph = bs.phase_breakdown(eng, wl, lr, dev, 8)
    # the same phases, each replayed alone as a CUDA graph (no host launch gaps; world 1: every phase is idempotent)
    st = torch.cuda.Stream(dev)
    calls = [("prepare", lambda: eng.prepare(0, wl.ids[0])), ("owner gather", lambda: eng.lookup(0, wl.out[0], phases=(1, 1))),
             ("expand", lambda: eng.lookup(0, wl.out[0], phases=(3, 3))),
             ("local MergeAdd", lambda: eng.update(0, wl.dout[0], lr, phases=(0, 0)))]
    with torch.cuda.stream(st):
        eng.prepare(0, wl.ids[0]); eng.lookup(0, wl.out[0]); eng.update(0, wl.dout[0], lr)
        st.synchronize()
        for name, fn in calls:
            fn(); st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st, capture_error_mode="thread_local"):
                fn()
            for _ in range(3):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(30):
                g.replay()
            e1.record(st)
            st.synchronize()
            print("   graph replay  %8.1f us  %s" % (e0.elapsed_time(e1) / 30 * 1e3, name))
    print("world 1, %d ids: sum %.1f us" % (n, sum(list(ph.values())[1:])))
    for k, x in ph.items():
        print("   %8.1f us  %s" % (x, k))
    eng.destroy()
dist.destroy_process_group()
