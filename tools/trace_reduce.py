#!/usr/bin/env python
"""Developer tool: where the time of one k_chunk_reduce launch goes.  Builds a -DB200EMB_TRACE variant of
the library (tools/_trace/libb200emb_trace.so — `python tools/trace_reduce.py --build` here, the .so
travels to the GPU box), runs the fused update of one config with a ready plan and prints per-phase
statistics of the per-CTA %globaltimer stamps:
  0 start | 1 keys in smem | 2 main loop done (group 0) | 3 flags barrier | 4 combine done | 5 arrival done | 6 chain closed
"""
import argparse
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT_DIR = os.path.join(ROOT, "tools", "_trace")
LIB = os.path.join(OUT_DIR, "libb200emb_trace.so")


def build():
    from __graft_entry__ import load_package
    pkg = load_package()
    import importlib
    init = importlib.import_module("paddle1_b200")
    os.makedirs(OUT_DIR, exist_ok=True)
    csrc = os.path.join(ROOT, "paddle-1_b200", "csrc")
    srcs = [os.path.join(csrc, s) for s in init.KERNEL_SOURCES if "f64" not in s]
    # the f64 instantiation is needed for linking capi.cu: keep it, without the trace macro it is unchanged
    srcs.append(os.path.join(csrc, "segreduce_f64.cu"))
    cmd = [init._nvcc()] + init.NVCC_FLAGS + ["-DB200EMB_TRACE", "-o", LIB] + srcs
    subprocess.check_call(cmd)
    print("built", LIB)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build", action="store_true")
    ap.add_argument("--config", default="C2_ctr_sgd")
    ap.add_argument("--launches", type=int, default=5)
    args = ap.parse_args()
    if args.build:
        build()
        return
    import numpy as np
    import torch
    from __graft_entry__ import load_package
    pkg = load_package()
    capi, synth = pkg.capi, pkg.synth
    capi._LIB_PATH = LIB
    capi._lib = None
    lib = capi.lib()
    v, d, n, a, opt = synth.CONFIGS[args.config]
    dev = torch.device("cuda:0")
    z = synth.ZipfIds(v, a, seed=0)
    nb = 6
    ids = [torch.from_numpy(z.sample(n)).to(dev) for _ in range(nb)]
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    table = torch.empty((v, d), device=dev).uniform_(-0.05, 0.05, generator=g)
    grads = [torch.randn((n, d), device=dev, generator=g) * 0.01 for _ in range(nb)]
    outs = [torch.empty((n, d), device=dev) for _ in range(nb)]
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(nb)]
    lr = torch.tensor([0.001], device=dev)
    for b in range(nb):
        capi.plan_rows(ids[b], v, ws[b])
    torch.cuda.synchronize()
    rows = 8192
    buf = np.zeros((rows, 8), dtype=np.uint64)
    for it in range(args.launches):
        b = it % nb
        # evict: stream the other buffers through the gather so this batch is cold in L2
        for o in range(nb):
            if o != b:
                capi.lookup_table(table, ids[o], -1, outs[o])
        lib.b200emb_debug_trace_clear()
        torch.cuda.synchronize()
        capi.sparse_sgd(table, ids[b], grads[b], lr, ws[b], capi.PLAN_READY)
        torch.cuda.synchronize()
        lib.b200emb_debug_trace(buf.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(buf.nbytes))
        t = buf.astype(np.int64)
        live = t[:, 0] > 0
        t = t[live]
        t0 = t[:, 0].min()
        end = np.where(t[:, 6] > 0, t[:, 6], np.where(t[:, 5] > 0, t[:, 5], t[:, 4]))
        print("launch %d: %d CTAs, kernel span %.2f us (first start -> last end)" % (it, len(t), (end.max() - t0) / 1e3))
        print("  CTA start spread: %.2f us" % ((t[:, 0].max() - t0) / 1e3))
        for name, a0, a1 in (("keys->smem", 0, 1), ("main loop (group 0)", 1, 2), ("flags barrier wait", 2, 3),
                             ("combine+tail consume", 3, 4)):
            dt = (t[:, a1] - t[:, a0]) / 1e3
            print("  %-24s median %.2f  p90 %.2f  max %.2f us" % (name, np.median(dt), np.percentile(dt, 90), dt.max()))
        arr = t[t[:, 5] > 0]
        if len(arr):
            dt = (arr[:, 5] - arr[:, 4]) / 1e3
            print("  %-24s median %.2f  p90 %.2f  max %.2f us  (%d CTAs)" % ("arrival (probe+fence+atom)", np.median(dt), np.percentile(dt, 90), dt.max(), len(arr)))
        cl = t[t[:, 6] > 0]
        if len(cl):
            dt = (cl[:, 6] - cl[:, 5]) / 1e3
            print("  %-24s median %.2f  p90 %.2f  max %.2f us  (%d CTAs)" % ("close_chain", np.median(dt), np.percentile(dt, 90), dt.max(), len(cl)))
        ids_live = np.nonzero(live)[0]
        pro = (t[:, 1] - t[:, 0]) / 1e3
        for j in np.argsort(-pro)[:6]:
            print("    slow prologue: CTA %d  keys->smem %.2f  main %.2f  tail %.2f  arrival? %s close? %s" % (
                ids_live[j], pro[j], (t[j, 2] - t[j, 1]) / 1e3, (t[j, 4] - t[j, 3]) / 1e3, t[j, 5] > 0, t[j, 6] > 0))
        late = (t[:, 2] - t0) / 1e3
        for j in np.argsort(-late)[:6]:
            print("    late main-loop end: CTA %d  end %.2f  keys->smem %.2f  main %.2f" % (
                ids_live[j], late[j], pro[j], (t[j, 2] - t[j, 1]) / 1e3))
        m_end = (t[:, 2] - t0) / 1e3
        print("  main-loop end after first start: p10 %.2f  median %.2f  p90 %.2f  max %.2f us" % (
            np.percentile(m_end, 10), np.median(m_end), np.percentile(m_end, 90), m_end.max()))
        print("  time of last main-loop end %.2f us, last combine end %.2f us, last end %.2f us after first start" % (
            (t[:, 2].max() - t0) / 1e3, (t[:, 4].max() - t0) / 1e3, (end.max() - t0) / 1e3))


if __name__ == "__main__":
    main()
