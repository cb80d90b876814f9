"""Developer tool: runs the large-batch sharded-path kernels once per iteration (for an ncu launch list)."""
import sys, torch, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load_package(); capi = pkg.capi
n, v, G, d = 1703936, 100_000_000, 8, 64
ids = torch.from_numpy(pkg.synth.ZipfIds(v, 1.1, seed=0).sample(n)).cuda()
ws = capi.alloc_workspace(n, d)
dout = torch.randn((n, d), device="cuda")
for it in range(3):
    uk, nu, b, eb, loc = capi.plan_unique_rows(ids, v, G, ws)
    u = int(nu.item())
    uv = torch.randn((u, d), device="cuda")
    out = capi.expand_unique(uv, n, capi.plan_height(v, G), ws)
    r, m, _ = capi.merge_add(ids, dout, capi.plan_height(v, G), workspace=ws, flags=capi.PLAN_READY)
    shard = torch.zeros((v // G, d), device="cuda")
    lr = torch.tensor([0.01], device="cuda")
    capi.sparse_sgd(shard, loc[:u], m[:u], lr)
torch.cuda.synchronize()
print("ok", u)
