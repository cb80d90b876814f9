#!/usr/bin/env python
"""Developer tool: per-phase device times (gather / sort plan / fused update with PLAN_READY) of one
BASELINE.json config on cuda:0, each phase replayed as CUDA graphs over rotating batch buffers larger than
L2.  Prints one JSON line.  Kernel tuning switches (B200EMB_*) are read from the environment by the
library, so sweeps run this script once per setting.

  python tools/phase_bench.py C2_ctr_sgd [--reps 200] [--det] [--tag text]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--det", action="store_true", help="also time the bit-exact order mode")
    ap.add_argument("--tag", default="")
    ap.add_argument("--det-only", action="store_true", help="three eager bit-exact-mode updates and nothing else (for ncu)")
    args = ap.parse_args()
    pkg = load_package()
    capi, synth = pkg.capi, pkg.synth
    capi.lib()
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    v, d, n, a, opt = synth.CONFIGS[args.config]
    dev = torch.device("cuda:0")
    per_buf = n * (8 + 2 * 4 * d)
    nb = max(2, min(8, int(np.ceil(400e6 / per_buf))))
    z = synth.ZipfIds(v, a, seed=0)
    ids = [torch.from_numpy(z.sample(n)).to(dev) for _ in range(nb)]
    u = float(np.mean([int(torch.unique(i).numel()) for i in ids]))
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    table = torch.empty((v, d), device=dev)
    for lo in range(0, v, 1 << 22):
        table[lo:lo + (1 << 22)].uniform_(-0.05, 0.05, generator=g)
    grads = [torch.randn((n, d), device=dev, generator=g) * 0.01 for _ in range(nb)]
    outs = [torch.empty((n, d), device=dev) for _ in range(nb)]
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(nb)]
    status = capi.new_status(dev)
    lr = torch.tensor([0.001], device=dev)
    m1 = m2 = b1p = b2p = None
    if opt == "adam":
        m1, m2 = torch.zeros_like(table), torch.zeros_like(table)
        b1p, b2p = torch.tensor([0.9], device=dev), torch.tensor([0.999], device=dev)

    def update(b, flags):
        if opt == "adam":
            capi.sparse_adam(table, m1, m2, ids[b], grads[b], lr, b1p, b2p, workspace=ws[b], flags=flags, status=status)
        else:
            capi.sparse_sgd(table, ids[b], grads[b], lr, ws[b], flags, status)

    stream = torch.cuda.Stream(dev)
    if args.det_only:
        for b in range(3):
            update(b % nb, capi.DETERMINISTIC)
        torch.cuda.synchronize()
        print(json.dumps({"config": args.config, "det_only": True, "status_word": int(status.item())}))
        return

    def phase(fn, reps):
        with torch.cuda.stream(stream):
            for b in range(nb):
                fn(b)
            stream.synchronize()
            gs = []
            for b in range(nb):
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=stream):
                    fn(b)
                gs.append(gr)
            for b in range(nb):
                gs[b].replay()
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for i in range(reps):
                gs[i % nb].replay()
            e1.record(stream)
            stream.synchronize()
        return e0.elapsed_time(e1) / reps * 1e-3

    reps = args.reps
    by = synth.algorithmic_bytes(n, d, u, opt)
    t_gather = phase(lambda b: capi.lookup_table(table, ids[b], -1, outs[b], status), reps)
    t_plan = phase(lambda b: capi.plan_rows(ids[b], v, ws[b], status), reps)
    for b in range(nb):
        capi.plan_rows(ids[b], v, ws[b], status)
    t_upd = phase(lambda b: update(b, capi.PLAN_READY), reps)
    t_serial = phase(lambda b: (capi.lookup_table(table, ids[b], -1, outs[b], status), update(b, 0)), reps)
    res = {"config": args.config, "tag": args.tag, "table": [v, d], "ids_per_step": n, "unique": u, "optimizer": opt,
           "buffers": nb, "gather_us": t_gather * 1e6, "gather_frac": by["fwd"] / t_gather / 1e9 / peak,
           "sort_plan_us": t_plan * 1e6, "update_us": t_upd * 1e6, "update_frac": by["upd"] / t_upd / 1e9 / peak,
           "serial_step_us": t_serial * 1e6, "serial_step_frac": by["step"] / t_serial / 1e9 / peak,
           "overlapped_bound_frac": by["step"] / (t_gather + t_upd) / 1e9 / peak}
    if args.det:
        t_det = phase(lambda b: update(b, capi.DETERMINISTIC), max(3, reps // 20))
        res["det_update_incl_sort_us"] = t_det * 1e6
    res["status_word"] = int(status.item())
    print(json.dumps(res))


if __name__ == "__main__":
    main()
