#!/usr/bin/env python
"""Secondary measurements for BASELINE.md §5: device-resident step time of the large single-GPU configs
(C3 lazy Adam 10M x 128, C5 duplicate-heavy SGD 50M x 256) — forward gather + fused sort/reduce/update —
with the fraction of the measured HBM roofline.  Not the contract bench (that is bench.py on configs[1])."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


def main():
    pkg = load_package()
    capi, synth = pkg.capi, pkg.synth
    capi.lib()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for name in sys.argv[1:] or ["C3_adam", "C5_dup_heavy"]:
        v, d, n, a, opt = synth.CONFIGS[name]
        z = synth.ZipfIds(v, a, seed=0)
        nb = 3
        ids = [torch.from_numpy(z.sample(n)).cuda() for _ in range(nb)]
        u = float(np.mean([len(torch.unique(i)) for i in ids]))
        g = torch.Generator(device="cuda")
        g.manual_seed(1)
        table = torch.empty((v, d), device="cuda")
        for lo in range(0, v, 1 << 22):
            table[lo:lo + (1 << 22)].uniform_(-0.05, 0.05, generator=g)
        grads = [torch.randn((n, d), device="cuda", generator=g) * 0.01 for _ in range(nb)]
        outs = [torch.empty((n, d), device="cuda") for _ in range(nb)]
        ws = capi.alloc_workspace(n, d)
        lr = torch.tensor([0.001], device="cuda")
        if opt == "adam":
            m1, m2 = torch.zeros_like(table), torch.zeros_like(table)
            b1p, b2p = torch.tensor([0.9], device="cuda"), torch.tensor([0.999], device="cuda")

        def step(b):
            capi.lookup_table(table, ids[b], -1, outs[b])
            if opt == "adam":
                capi.sparse_adam(table, m1, m2, ids[b], grads[b], lr, b1p, b2p, workspace=ws)
            else:
                capi.sparse_sgd(table, ids[b], grads[b], lr, workspace=ws)

        def phase(fn, reps):
            for b in range(nb):
                fn(b)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(reps):
                fn(i % nb)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps * 1e-3

        reps = 20
        t_step = phase(step, reps)
        t_gather = phase(lambda b: capi.lookup_table(table, ids[b], -1, outs[b]), reps)
        t_plan = phase(lambda b: capi.plan_rows(ids[b], v, ws), reps)
        by = synth.algorithmic_bytes(n, d, u, opt)
        print(json.dumps({"config": name, "table": [v, d], "ids_per_step": n, "unique": u, "optimizer": opt,
                          "step_us": t_step * 1e6, "ids_per_s": n / t_step,
                          "step_frac_of_%g_GBs" % peak: by["step"] / t_step / 1e9 / peak,
                          "gather_us": t_gather * 1e6, "gather_frac": by["fwd"] / t_gather / 1e9 / peak,
                          "sort_plan_us": t_plan * 1e6,
                          "update_incl_sort_us": (t_step - t_gather) * 1e6,
                          "update_frac": by["upd"] / (t_step - t_gather) / 1e9 / peak}))
        del table, grads, outs
        if opt == "adam":
            del m1, m2
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
