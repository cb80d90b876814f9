#!/usr/bin/env python
"""bench.py — sparse-embedding hot path (lookup_table -> SelectedRows grad -> fused MergeAdd+SGD).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU kernels on host cores

One "step" = one pass of the hot path over one synthetic batch:
  N=1 : BASELINE.json configs[1] — 1M x 64 fp32 table, 4096 x 26 Zipf(1.1) ids, sparse SGD:
        forward gather of the batch's rows, then the sort-by-id + duplicate-row reduction fused
        with the in-place SGD row update (the SelectedRows gradient aliases ids/dOut: zero copy).
  N>1 : configs[3] — 100M x 64 table row-sharded over the N GPUs (owner = id % N), the same
        per-GPU batch; id all-to-all -> gather -> row all-to-all; grad all-to-all -> fused update.
Prints ONE JSON line (rank 0).  metric = lookups+updates/s = ids processed per second, whole job.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "embedding lookups+updates/sec"
UNIT = "ids/s"
N_ROTATE = 8  # distinct input/output batch buffers cycled through (together > L2)


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    _NAMES = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
              0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
              0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index=0, period=0.004):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self._NAMES.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def load_pkg():
    from __graft_entry__ import load_package
    return load_package()


# --------------------------------------------------------------------------- CPU arms
def cpu_step_fn():
    """(fn(table, ids, dout, lr) -> out, kind).  kind 'reference' = oracle/_ref (the reference's own CPU
    kernels compiled from /root/reference at build time) when it was built, else the oracle 'port'."""
    import oracle
    ref = None
    try:
        from oracle import ref_cpu
        if ref_cpu.available():
            ref = ref_cpu
    except Exception:
        ref = None
    impl = ref if ref is not None else oracle

    def fn(table, ids, dout, lr):
        out = impl.lookup_table(table, ids)
        rows, value = impl.lookup_table_grad_sparse(ids, dout)
        impl.sgd_sparse(table, rows, value, lr)
        return out
    return fn, ("reference" if ref is not None else "port")


def time_cpu(cfg, budget_s=None, steps=None, warmup=1):
    synth = load_pkg().synth
    v, d, n, a, _ = cfg
    table = synth.table_values(v, d)
    lr = np.array([0.01], np.float32)
    fn, kind = cpu_step_fn()
    z = synth.ZipfIds(v, a, seed=0)
    batches = [(z.sample(n), synth.grad_values(n, d, seed=4321 + i)) for i in range(2)]
    for w in range(max(1, warmup)):
        fn(table, batches[w % 2][0], batches[w % 2][1], lr)
    done, t0 = 0, time.perf_counter()
    while True:
        ids, dout = batches[done % 2]
        fn(table, ids, dout, lr)
        done += 1
        el = time.perf_counter() - t0
        if (steps is not None and done >= steps) or (steps is None and el >= budget_s):
            break
    el = time.perf_counter() - t0
    return {"value": n * done / el, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": "%d full steps of the workload (%d ids each: fwd gather + SelectedRows grad + "
                      "sparse SGD), one thread as the reference runs these loops, %.2f s" % (done, n, el),
            "ms_per_step": el / done * 1e3}


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    pkg = load_pkg()
    name = "C2_ctr_sgd" if args.gpus == 1 else "C4_sharded_sgd"
    v, d, n, a, opt = pkg.synth.CONFIGS[name]
    if args.gpus > 1:
        # the reference replicates the table on every device; its CPU kernels process the
        # same N ids per step against one table.  100M x 64 does not fit host RAM comfortably: the CPU
        # arm uses the per-shard table (V/gpus rows) like one owner does.
        v = v // args.gpus
    steps = max(1, min(args.steps, 200))
    res = time_cpu((v, d, n, a, opt), steps=steps, warmup=min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": res["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(args.gpus), "table": [v, d], "ids_per_step": n,
                   "optimizer": opt, "note": "reference CPU kernels (%s), single thread" % res["kind"]},
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_name(gpus):
    if gpus == 1:
        return "configs[1]: synthetic CTR 1Mx64 fp32 table, batch 4096x26 Zipf(1.1) ids, sparse SGD, 1xB200"
    return ("configs[3]: synthetic 100Mx64 fp32 table row-sharded over %d B200 (id %% G), 4096x26 Zipf(1.1) ids "
            "per GPU, id/row/grad all-to-all, sparse SGD" % gpus)


# --------------------------------------------------------------------------- single GPU
def run_single(args):
    import torch
    pkg = load_pkg()
    capi, synth = pkg.capi, pkg.synth
    capi.lib()  # raises if the CUDA library is missing: no fallback
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (use --impl reference for the CPU arm)")
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    v, d, n, a, opt = synth.CONFIGS["C2_ctr_sgd"]
    peak, peak_src = measured_peak_gbs()

    table = torch.from_numpy(synth.table_values(v, d)).to(dev)
    lr = torch.tensor([0.01], dtype=torch.float32, device=dev)
    z = synth.ZipfIds(v, a, seed=0)
    ids_h, dout_h, uniq = [], [], []
    for b in range(N_ROTATE):
        i = z.sample(n)
        ids_h.append(torch.from_numpy(i).pin_memory())
        dout_h.append(torch.from_numpy(synth.grad_values(n, d, seed=4321 + b)).pin_memory())
        uniq.append(len(np.unique(i)))
    u_mean = float(np.mean(uniq))
    ids = [t.to(dev) for t in ids_h]
    dout = [t.to(dev) for t in dout_h]
    out = [torch.empty((n, d), dtype=torch.float32, device=dev) for _ in range(N_ROTATE)]
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(N_ROTATE)]
    status = capi.new_status(dev)
    bytes_ = synth.algorithmic_bytes(n, d, u_mean, opt)

    # the critical path (gather -> reduce/update) gets the high-priority stream; the sort plans of the
    # following batches fill in behind it on normal-priority streams
    main = torch.cuda.Stream(dev, priority=-1)
    side = torch.cuda.Stream(dev, priority=0)

    def step(b):
        """One step on batch b.  The sort plan of the NEXT batch (ids only, like a data-loader prefetch)
        runs on the side stream while this batch's gather + fused reduce/update run on the main one;
        every step still launches exactly one plan, one gather, one update."""
        nb = (b + 1) % N_ROTATE
        side.wait_stream(main)
        with torch.cuda.stream(side):
            capi.plan_rows(ids[nb], v, ws[nb], status)
        capi.lookup_table(table, ids[b], -1, out[b], status)
        # lookup_table_grad (sparse): the SelectedRows gradient aliases ids / dOut — zero copy
        capi.sparse_sgd(table, ids[b], dout[b], lr, ws[b], capi.PLAN_READY, status)
        main.wait_stream(side)

    # gather, sort plan, chunk reduce (fast-mode SGD finds the run heads itself: no head scan)
    kernels_per_step = 1 + capi.plan_kernel_count(n, v) + 1

    graphs = []
    with torch.cuda.stream(main):
        for b in range(N_ROTATE):   # every workspace holds its batch's plan before the first step
            capi.plan_rows(ids[b], v, ws[b], status)
        for b in range(N_ROTATE):   # warm every path once outside capture
            step(b)
        main.synchronize()
        for b in range(N_ROTATE):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=main):
                step(b)
            graphs.append(g)
        # one graph holding a whole rotation (N_ROTATE consecutive steps): amortises the per-graph launch gap
        rotation = torch.cuda.CUDAGraph()
        side2 = torch.cuda.Stream(dev, priority=0)
        with torch.cuda.graph(rotation, stream=main):
            # Inside a rotation the plan of batch b+1 only has to be ready when update(b+1) starts, so the
            # plans run as two alternating side chains that overlap gather AND update of the main stream.
            sides = [side, side2]
            for st in sides:
                st.wait_stream(main)
            plan_done, upd0_done = {}, None
            for b in range(N_ROTATE):
                nb = (b + 1) % N_ROTATE
                st = sides[nb % 2]
                with torch.cuda.stream(st):
                    if nb == 0:
                        st.wait_event(upd0_done)          # ws[0] is still read by this rotation's update(0)
                    capi.plan_rows(ids[nb], v, ws[nb], status)
                    plan_done[nb] = torch.cuda.Event()
                    plan_done[nb].record(st)
                capi.lookup_table(table, ids[b], -1, out[b], status)
                if b in plan_done:
                    main.wait_event(plan_done[b])
                capi.sparse_sgd(table, ids[b], dout[b], lr, ws[b], capi.PLAN_READY, status)
                if b == 0:
                    upd0_done = torch.cuda.Event()
                    upd0_done.record(main)
            for st in sides:
                main.wait_stream(st)
    torch.cuda.synchronize()

    def run_steps(k, start=0):
        """k consecutive steps starting at batch `start % N_ROTATE`: whole rotations as one graph launch,
        the ragged head/tail as single-step graphs."""
        i = 0
        while i < k and (start + i) % N_ROTATE != 0:
            graphs[(start + i) % N_ROTATE].replay()
            i += 1
        while k - i >= N_ROTATE:
            rotation.replay()
            i += N_ROTATE
        while i < k:
            graphs[(start + i) % N_ROTATE].replay()
            i += 1

    with torch.cuda.stream(main):
        run_steps(args.warmup)
        torch.cuda.synchronize()
        sampler = ClockSampler(0)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        run_steps(args.steps, args.warmup)
        e1.record(main)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        # keep the GPU under the same load a little longer so the clock sampler sees it
        t_end = time.perf_counter() + 0.25
        while time.perf_counter() < t_end:
            run_steps(50)
            torch.cuda.synchronize()
        clocks = sampler.stop()
    ms_per_step = ms / args.steps
    value = n * args.steps / (ms * 1e-3)

    # ---- per-phase timings (CUDA events on the launching stream, same rotating buffers) ----------
    def time_phase(fn, reps):
        with torch.cuda.stream(main):
            for b in range(N_ROTATE):
                fn(b)
            main.synchronize()
            gs = []
            for b in range(N_ROTATE):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=main):
                    fn(b)
                gs.append(g)
            for i in range(N_ROTATE):
                gs[i].replay()
            main.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(main)
            for i in range(reps):
                gs[i % N_ROTATE].replay()
            a1.record(main)
            main.synchronize()
            return a0.elapsed_time(a1) / reps * 1e-3

    reps = max(50, min(args.steps, 400))
    t_gather = time_phase(lambda b: capi.lookup_table(table, ids[b], -1, out[b], status), reps)
    t_plan = time_phase(lambda b: capi.plan_rows(ids[b], v, ws[b], status), reps)
    for b in range(N_ROTATE):
        capi.plan_rows(ids[b], v, ws[b], status)
    t_update = time_phase(lambda b: capi.sparse_sgd(table, ids[b], dout[b], lr, ws[b], capi.PLAN_READY, status), reps)
    phases = {
        "gather_us": t_gather * 1e6, "sort_plan_us": t_plan * 1e6, "reduce_update_us": t_update * 1e6,
        "gather_GBs": bytes_["fwd"] / t_gather / 1e9, "reduce_update_GBs": bytes_["upd"] / t_update / 1e9,
    }
    if t_gather >= t_update:
        dom = {"kernel": "k_gather (lookup_table fwd)", "bytes": bytes_["fwd"], "t": t_gather}
    else:
        dom = {"kernel": "k_chunk_reduce (fused MergeAdd+SGD)", "bytes": bytes_["upd"], "t": t_update}
    achieved = dom["bytes"] / dom["t"] / 1e9
    # DRAM traffic of the same kernel(s) from the committed `ncu --set full` capture (profiles/), per launch
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r1_ncu_full_summary.json")) as f:
            prof = json.load(f)
        keys = ["k_gather"] if dom["kernel"].startswith("k_gather") else ["k_chunk_reduce"]
        traffic = sum(v["dram_traffic_bytes_per_launch"] for k, v in prof.items() if any(k.startswith(q) for q in keys))
    except Exception:
        traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": dom["kernel"], "algorithmic_bytes_per_launch": dom["bytes"],
                "peak_source": peak_src,
                "step": {"algorithmic_bytes": bytes_["step"], "achieved": bytes_["step"] / (ms_per_step * 1e-3) / 1e9,
                         "frac": bytes_["step"] / (ms_per_step * 1e-3) / 1e9 / peak}}

    # ---- end to end: host buffers, H2D of ids+dOut and D2H of Out inside the timed region ---------
    e2e = run_e2e(torch, capi, table, lr, ids_h, dout_h, n, d, v, status, max(20, min(args.steps, 200)))

    st = int(status.item())
    cpu = time_cpu(synth.CONFIGS["C2_ctr_sgd"], budget_s=10.0)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(1), "table": [v, d], "ids_per_step": n,
                   "unique_ids_per_step": u_mean, "optimizer": opt, "mode": "fast (fixed-tree reduction)",
                   "l2": "inputs rotate over %d batch buffers (%.0f MB ids+dOut+Out) > 126 MB L2; table 256 MB"
                         % (N_ROTATE, N_ROTATE * (n * 8 + 2 * n * d * 4) / 1e6),
                   "launch": "CUDA graphs (one per rotation of %d steps, single-step graphs for the remainder);" % N_ROTATE + " the sort plan of batch i+1 runs on a side stream during step i (ids-only prefetch), gather+update of batch i on the main stream"},
        "roofline": roofline, "phases": phases, "cpu_baseline": {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": e2e, "gpu_launches": kernels_per_step * args.steps, "clocks": clocks, "status_word": st,
    }
    print(json.dumps(line))


def run_e2e(torch, capi, table, lr, ids_h, dout_h, n, d, v, status, steps):
    """Public-API step on HOST buffers: per step H2D(ids, dOut) from pinned memory, the hot path, D2H(Out).
    Three streams (H2D / compute / D2H) with double-buffered device staging so copies of neighbouring
    steps overlap the kernels, as a feed/fetch pipeline would."""
    dev = table.device
    nbuf = 2
    ids_d = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(nbuf)]
    dout_d = [torch.empty((n, d), dtype=torch.float32, device=dev) for _ in range(nbuf)]
    out_d = [torch.empty((n, d), dtype=torch.float32, device=dev) for _ in range(nbuf)]
    out_h = [torch.empty((n, d), dtype=torch.float32).pin_memory() for _ in range(nbuf)]
    ws = [capi.alloc_workspace(n, d, 4, dev) for _ in range(nbuf)]
    s_in, s_cmp, s_out = (torch.cuda.Stream(dev) for _ in range(3))
    ev_in = [torch.cuda.Event() for _ in range(nbuf)]
    ev_cmp = [torch.cuda.Event() for _ in range(nbuf)]
    ev_out = [torch.cuda.Event() for _ in range(nbuf)]

    def one(i):
        b = i % nbuf
        src = i % len(ids_h)
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_cmp[b])      # staging buffers free once the step that used them computed
            ids_d[b].copy_(ids_h[src], non_blocking=True)
            dout_d[b].copy_(dout_h[src], non_blocking=True)
            ev_in[b].record(s_in)
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(ev_in[b])
            s_cmp.wait_event(ev_out[b])     # out_d[b] drained by the previous user
            capi.lookup_table(table, ids_d[b], -1, out_d[b], status)
            capi.sparse_sgd(table, ids_d[b], dout_d[b], lr, ws[b], 0, status)
            ev_cmp[b].record(s_cmp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_cmp[b])
            out_h[b].copy_(out_d[b], non_blocking=True)
            ev_out[b].record(s_out)

    for i in range(4):
        one(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    return {"value": n * steps / el, "unit": UNIT, "h2d_bytes_per_step": n * 8 + n * d * 4,
            "d2h_bytes_per_step": n * d * 4, "ms_per_step": el / steps * 1e3, "steps": steps,
            "api": "capi.lookup_table + capi.sparse_sgd (C ABI) on pinned host batches, 3-stream feed/fetch pipeline"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=40)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ids-per-gpu", type=int, default=0, dest="ids_per_gpu",
                    help="sharded runs: ids per GPU per step (default 16 x 4096 x 26)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference_arm(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        from importlib import import_module
        sys.path.insert(0, ROOT)
        load_pkg()
        import_module("paddle1_b200.sharded_bench").run(args)
        return
    run_single(args)


if __name__ == "__main__":
    main()
