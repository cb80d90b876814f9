"""Developer tool: host-side time of one sharded step on one GPU (world 1), inline vs prefetched prepare."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load_package(); capi = pkg.capi
n, v, d = 1703936, 12_500_000, 64
dev = torch.device("cuda")
tab = pkg.sharded.ShardedTable(v, d, pkg.sharded.CudaOps(capi), dev)
ids = [torch.from_numpy(pkg.synth.ZipfIds(v, 1.1, seed=s).sample(n)).cuda() for s in range(4)]
dout = torch.randn((n, d), device=dev)
ws = [capi.alloc_workspace(n, d) for _ in range(4)]
ows = capi.alloc_workspace_upto(2 * n, d, 4, dev)
lr = torch.tensor([0.01], device=dev)
side = torch.cuda.Stream()
for mode in ("inline", "prefetch", "inline"):
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); tp = tl = tu = 0.0
        prep = tab.prepare(ids[0], ws[0], stream=side) if mode == "prefetch" else None
        for i in range(20):
            b = i % 4
            a = time.perf_counter()
            nxt = tab.prepare(ids[(i + 1) % 4], ws[(i + 1) % 4], stream=side) if (mode == "prefetch" and i < 19) else None
            c = time.perf_counter()
            out, ctx = tab.lookup(ids[b], ws[b], prep=prep)
            e = time.perf_counter()
            tab.update(ctx, dout, lr, owner_ws=ows)
            f = time.perf_counter()
            tp += c - a; tl += e - c; tu += f - e
            prep = nxt
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        print(mode, "ms/step %.3f  host: prepare %.3f lookup %.3f update %.3f" % ((t1 - t0) / 20 * 1e3, tp / 20 * 1e3, tl / 20 * 1e3, tu / 20 * 1e3))
