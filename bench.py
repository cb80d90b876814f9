#!/usr/bin/env python
"""bench.py — sparse-embedding hot path (lookup_table -> SelectedRows grad -> fused MergeAdd+SGD).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU kernels on host cores

One "step" = one pass of the hot path over one synthetic batch:
  N=1 : BASELINE.json configs[1] — 1M x 64 fp32 table, 4096 x 26 Zipf(1.1) ids, sparse SGD:
        forward gather of the batch's rows, then the sort-by-id + duplicate-row reduction fused
        with the in-place SGD row update (the SelectedRows gradient aliases ids/dOut: zero copy).
  N>1 : configs[3] — 100M x 64 table row-sharded over the N GPUs (owner = id % N), the same
        per-GPU batch, through the row-sharded step engine (bench_sharded.py): ids -> owners, owner gather ->
        requesters, local MergeAdd -> owners, fused owner update; plus a batch sweep and a parity epilogue.
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


def _table_in_place(height, d, seed):
    """uniform(-0.05, 0.05) table filled chunk by chunk (no full-size temporaries: N of these are alive at once)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    t = np.empty((height, d), np.float32)
    for lo in range(0, height, 1 << 20):
        hi = min(lo + (1 << 20), height)
        rng.random((hi - lo, d), dtype=np.float32, out=t[lo:hi])
        t[lo:hi] *= np.float32(0.1)
        t[lo:hi] -= np.float32(0.05)
    return t


def time_cpu_job(cfg, replicas, steps, warmup=1):
    """The reference's data-parallel job on the host cores: one replica per thread (its CPU kernels are
    single-threaded loops; ctypes releases the GIL).  Per step every replica looks its own ids up in its table and
    builds its SelectedRows gradient; the gradients of all replicas are gathered (concatenated in replica order —
    reduce_op_handle.cc:153-209 / the all-gather of the AllReduce strategy) and every replica applies the gathered
    gradient to its table.  value = replicas * ids per replica per step / wall time."""
    import threading
    synth = load_pkg().synth
    v, d, n, a, _ = cfg
    fn_impl, kind = _ref_module()
    lr = np.array([0.01], np.float32)
    tables = [_table_in_place(v, d, 1234 + r) for r in range(replicas)]
    z = [synth.ZipfIds(v, a, seed=r) for r in range(replicas)]
    batches = [[(z[r].sample(n), synth.grad_values(n, d, seed=4321 + 2 * r + i)) for i in range(2)] for r in range(replicas)]
    total = warmup + steps
    grads = [None] * replicas
    gate = threading.Barrier(replicas + 1)
    sync = threading.Barrier(replicas)

    def replica(r):
        for it in range(total):
            if it == warmup:
                gate.wait()
            ids, dout = batches[r][it % 2]
            fn_impl.lookup_table(tables[r], ids)
            grads[r] = fn_impl.lookup_table_grad_sparse(ids, dout)
            sync.wait()                                   # every replica's gradient is there
            rows = np.concatenate([g[0] for g in grads])
            value = np.concatenate([g[1] for g in grads])
            fn_impl.sgd_sparse(tables[r], rows, value, lr)
            sync.wait()                                   # nobody overwrites a gradient another replica still reads
        gate.wait()

    th = [threading.Thread(target=replica, args=(r,)) for r in range(replicas)]
    for t in th:
        t.start()
    gate.wait()
    t0 = time.perf_counter()
    gate.wait()
    el = time.perf_counter() - t0
    for t in th:
        t.join()
    return {"value": replicas * n * steps / el, "unit": UNIT, "cores": replicas, "kind": kind,
            "sample": "%d full steps of the %d-replica job (%d ids per replica per step: fwd gather + SelectedRows grad, "
                      "gather of all replicas' gradients, sparse SGD of the gathered %d rows on every replica), one thread "
                      "per replica, %.2f s" % (steps, replicas, n, replicas * n, el),
            "ms_per_step": el / steps * 1e3}


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
    if args.gpus > 1:
        steps = min(steps, 40)
        res = time_cpu_job((v, d, n, a, opt), args.gpus, steps=steps, warmup=min(args.warmup, 2))
        note = ("reference CPU kernels (%s): the %d-replica data-parallel job, one host thread per replica; every replica "
                "holds a table of V/G rows (G full 100M-row replicas do not fit host RAM), looks up its own %d ids and "
                "applies the gathered gradient of all replicas, as the reference's multi-device graph does"
                % (res["kind"], args.gpus, n))
    else:
        res = time_cpu((v, d, n, a, opt), steps=steps, warmup=min(args.warmup, 3))
        note = "reference CPU kernels (%s), single thread" % res["kind"]
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": res["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(args.gpus), "table": [v, d], "ids_per_step": n * args.gpus,
                   "ids_per_replica_per_step": n, "optimizer": opt, "note": note},
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
def _ref_module():
    """The CPU checker: oracle/_ref (the reference's own kernels) when built, else the oracle restatement."""
    import oracle
    try:
        from oracle import ref_cpu
        if ref_cpu.available():
            return ref_cpu, "reference"
    except Exception:
        pass
    return oracle, "port"


def parity_c2(torch, capi, synth, dev):
    """The north_star bars on configs[1] at full size, against the reference CPU kernels on the same inputs:
    gathered rows bit-exact; bit-exact mode bit-exact; fast mode — the literal metric (fraction and max of
    |fast - ref_fp32| / |ref_fp32| > 1e-6 over the updated table) beside the reordering bound vs fp64."""
    import oracle
    ref, kind = _ref_module()
    v, d, n, a, _ = synth.CONFIGS["C2_ctr_sgd"]
    table0 = synth.table_values(v, d)
    ids = synth.ZipfIds(v, a, seed=7).sample(n)
    grad = synth.grad_values(n, d, seed=77)
    lr = np.array([0.01], np.float32)
    want = table0.copy()
    want_out = ref.lookup_table(want, ids)
    ref.sgd_sparse(want, ids, grad, lr)
    d_ids, d_grad, d_lr = torch.from_numpy(ids).to(dev), torch.from_numpy(grad).to(dev), torch.from_numpy(lr).to(dev)
    res = {"checker": "oracle/_ref (reference CPU kernels, unmodified sources)" if kind == "reference" else "oracle (CPU restatement)",
           "config": "configs[1] full size, one step, seed 7"}
    t = torch.from_numpy(table0).to(dev)
    out = capi.lookup_table(t, d_ids)
    res["gather_bit_exact"] = bool(np.array_equal(out.cpu().numpy().view(np.uint32), want_out.view(np.uint32)))
    capi.sparse_sgd(t, d_ids, d_grad, d_lr, flags=capi.DETERMINISTIC)
    res["det_bit_exact"] = bool(np.array_equal(t.cpu().numpy().view(np.uint32), want.view(np.uint32)))
    t = torch.from_numpy(table0).to(dev)
    capi.sparse_sgd(t, d_ids, d_grad, d_lr)
    fast = t.cpu().numpy()
    touched = np.unique(ids)
    rel = np.abs(fast[touched].astype(np.float64) - want[touched]) / np.maximum(np.abs(want[touched]), 1e-30)
    res["fast_frac_gt_1e-6_rel_vs_ref"] = float((rel > 1e-6).mean())
    res["fast_max_rel_vs_ref"] = float(rel.max())
    rows, merged = oracle.merge_add(ids, grad, truth=True)
    _, l1 = oracle.merge_add(ids, np.abs(grad), truth=True)
    truth = table0[rows].astype(np.float64) - 0.01 * merged
    bound = 1e-6 * (np.abs(table0[rows]) + 0.01 * l1)
    res["fast_within_reordering_bound_vs_fp64"] = bool((np.abs(fast[rows] - truth) <= bound + 1e-30).all())
    err_fast = np.abs(fast[rows] - truth).max()
    err_ref = np.abs(want[rows] - truth).max()
    res["max_abs_err_vs_fp64"] = {"fast": float(err_fast), "reference_order_fp32": float(err_ref)}
    res["untouched_rows_unchanged"] = bool(np.array_equal(np.delete(fast, touched, axis=0)[:50000], np.delete(table0, touched, axis=0)[:50000]))
    res["ok"] = bool(res["gather_bit_exact"] and res["det_bit_exact"] and res["fast_within_reordering_bound_vs_fp64"]
                     and res["untouched_rows_unchanged"])
    return res


def extra_config(torch, capi, synth, stepbench, name, dev, peak, steps):
    """One of the larger single-GPU configs (configs[2] lazy Adam, configs[4] duplicate-heavy SGD): device-timed
    step through the same pipeline, per-phase times, roofline fractions, and a shortened parity check (sampled
    rows incl. the hottest id, bit-exact mode vs the oracle, fast mode within the reordering bound)."""
    import oracle
    v, d, n, a, opt = synth.CONFIGS[name]
    per_buf = n * (8 + 2 * 4 * d)
    nb = max(2, min(8, int(np.ceil(300e6 / per_buf))))
    z = synth.ZipfIds(v, a, seed=0)
    ids_np = [z.sample(n) for _ in range(nb)]
    del z
    u_mean = float(np.mean([len(np.unique(i)) for i in ids_np]))
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    table = torch.empty((v, d), dtype=torch.float32, device=dev)
    for lo in range(0, v, 1 << 22):
        table[lo:lo + (1 << 22)].uniform_(-0.05, 0.05, generator=g)
    ids = [torch.from_numpy(i).to(dev) for i in ids_np]
    dout = [torch.randn((n, d), device=dev, generator=g) * 0.01 for _ in range(nb)]
    lr_v = 0.001 if opt == "adam" else 0.01
    # ---- parity first (one step on batch 0, sampled rows), on the pristine table
    uniq, counts = np.unique(ids_np[0], return_counts=True)
    rng = np.random.default_rng(5)
    rows_np = np.unique(np.concatenate([[uniq[np.argmax(counts)]], rng.choice(uniq, size=min(256, uniq.size), replace=False)]))
    rows_d = torch.from_numpy(rows_np).to(dev)
    pos = torch.nonzero(torch.isin(ids[0], rows_d)).reshape(-1)
    sub_ids = torch.searchsorted(rows_d, ids[0][pos]).cpu().numpy()
    sub_grad = dout[0][pos].cpu().numpy()
    before = table[rows_d].cpu().numpy()
    lr_t = torch.tensor([lr_v], device=dev)
    parity = {"sampled_rows": int(rows_np.size), "max_duplicates_of_one_id": int(counts.max())}
    ws = capi.alloc_workspace(n, d, 4, dev)
    if opt == "adam":
        b1p, b2p = torch.tensor([0.9], device=dev), torch.tensor([0.999], device=dev)
        want = [before.copy(), np.zeros_like(before), np.zeros_like(before)]
        oracle.adam_sparse(want[0], want[1], want[2], sub_ids, sub_grad, np.array([lr_v], np.float32),
                           np.array([0.9], np.float32), np.array([0.999], np.float32))
        m1, m2 = torch.zeros_like(table), torch.zeros_like(table)
        capi.sparse_adam(table, m1, m2, ids[0], dout[0], lr_t, b1p, b2p, workspace=ws, flags=capi.DETERMINISTIC)
        got = [t[rows_d].cpu().numpy() for t in (table, m1, m2)]
        parity["det_bit_exact"] = bool(all(np.array_equal(x.view(np.uint32), w.view(np.uint32)) for x, w in zip(got, want)))
        table[rows_d] = torch.from_numpy(before).to(dev)      # undo on the sampled rows, then the fast mode on them
        m1.zero_(); m2.zero_()
        # (rows outside the sample were stepped once in bit-exact mode: irrelevant for the timing that follows)
        capi.sparse_adam(table, m1, m2, ids[0], dout[0], lr_t, b1p, b2p, workspace=ws)
        fm1 = m1[rows_d].cpu().numpy()
        _, l1 = oracle.merge_add(sub_ids, np.abs(sub_grad), truth=True)
        parity["fast_within_bound"] = bool((np.abs(fm1.astype(np.float64) - want[1]) <= 1e-6 * (np.abs(want[1]) + 0.1 * l1) + 1e-12).all())
        del m1, m2
    else:
        want = before.copy()
        oracle.sgd_sparse(want, sub_ids, sub_grad, np.array([lr_v], np.float32))
        capi.sparse_sgd(table, ids[0], dout[0], lr_t, workspace=ws, flags=capi.DETERMINISTIC)
        parity["det_bit_exact"] = bool(np.array_equal(table[rows_d].cpu().numpy().view(np.uint32), want.view(np.uint32)))
        table[rows_d] = torch.from_numpy(before).to(dev)
        capi.sparse_sgd(table, ids[0], dout[0], lr_t, workspace=ws)
        _, l1 = oracle.merge_add(sub_ids, np.abs(sub_grad), truth=True)
        _, msum = oracle.merge_add(sub_ids, sub_grad, truth=True)
        truth = before.astype(np.float64) - lr_v * msum
        parity["fast_within_bound"] = bool((np.abs(table[rows_d].cpu().numpy() - truth) <= 1e-6 * (np.abs(before) + lr_v * l1)).all())
    out0 = capi.lookup_table(table, ids[0][:65536])
    parity["gather_bit_exact"] = bool(torch.equal(out0, table[ids[0][:65536]]))
    parity["ok"] = bool(parity["det_bit_exact"] and parity["fast_within_bound"] and parity["gather_bit_exact"])
    del ws, out0
    # ---- timing
    pipe = stepbench.SingleGpuPipeline(table, ids, dout, opt, lr_v)
    pipe.capture()
    ms = pipe.time_steps(steps, max(3, nb))
    ph = pipe.phase_times(max(6, steps // 2))
    by = synth.algorithmic_bytes(n, d, u_mean, opt)
    t_step = ms / steps * 1e-3
    res = {"config": name, "table": [v, d], "ids_per_step": n, "unique_ids_per_step": u_mean, "optimizer": opt,
           "ms_per_step": ms / steps, "value": n / t_step, "unit": UNIT, "steps": steps, "batch_buffers": nb,
           "phases": {"gather_us": ph["gather"] * 1e6, "sort_plan_us": ph["sort_plan"] * 1e6,
                      "reduce_update_us": ph["update"] * 1e6, "serial_step_us": ph["serial"] * 1e6},
           "roofline": {"bound": "hbm", "unit": "GB/s", "peak": peak,
                        "step": {"algorithmic_bytes": by["step"], "achieved": by["step"] / t_step / 1e9,
                                 "frac": by["step"] / t_step / 1e9 / peak},
                        "gather": {"algorithmic_bytes": by["fwd"], "frac": by["fwd"] / ph["gather"] / 1e9 / peak},
                        "reduce_update": {"algorithmic_bytes": by["upd"], "frac": by["upd"] / ph["update"] / 1e9 / peak}},
           "parity": parity, "status_word": int(pipe.status.item())}
    del pipe, table, ids, dout
    torch.cuda.empty_cache()
    return res


def e2e_opkernel(torch, pkg, table_np, ids_h, dout_h, n, d, v, steps, fused):
    """The same step through the drop-in boundary a Fluid program reaches: Operator.run of the registered
    lookup_table -> lookup_table_grad (is_sparse) -> sgd CUDA OpKernels (paddle-1_b200/fluid/, REGISTER_OP_CUDA_KERNEL)
    on host numpy batches: feed (H2D) ids and Out@GRAD, run the three ops, fetch Out (D2H), synchronously like
    Executor::Run does (framework/executor.cc:332-340)."""
    from importlib import import_module
    core = import_module("paddle1_b200.fluid_lite")
    core.set_fused(fused)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    scope.var("W").get_tensor().set(table_np, place)
    scope.var("LearningRate").get_tensor().set(np.array([0.01], np.float32), place)
    lookup = core.Operator("lookup_table", W="W", Ids="Ids", Out="Out", is_sparse=True)
    grad = core.Operator("lookup_table_grad", W="W", Ids="Ids", **{"Out@GRAD": "Out@GRAD", "W@GRAD": "W@GRAD"}, is_sparse=True)
    sgd = core.Operator("sgd", Param="W", Grad="W@GRAD", LearningRate="LearningRate", ParamOut="W")
    ids_np = [t.numpy().reshape(-1, 1) for t in ids_h]
    dout_np = [t.numpy() for t in dout_h]

    def one(i):
        b = i % len(ids_np)
        scope.var("Ids").get_tensor().set(ids_np[b], place)
        scope.var("Out@GRAD").get_tensor().set(dout_np[b], place)
        lookup.run(scope, place)
        grad.run(scope, place)
        sgd.run(scope, place)
        return np.array(scope.find_var("Out").get_tensor())
    for i in range(3):
        one(i)
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    el = time.perf_counter() - t0
    core.set_fused(False)
    return {"value": n * steps / el, "unit": UNIT, "ms_per_step": el / steps * 1e3, "steps": steps,
            "mode": "fused (FLAGS_b200emb_fused: gradient stays on the device, sort plan started by the forward op)"
                    if fused else "materialised (rows written into the SelectedRows like the reference)",
            "h2d_bytes_per_step": n * 8 + n * d * 4, "d2h_bytes_per_step": n * d * 4,
            "api": "fluid_lite.Operator('lookup_table' | 'lookup_table_grad' | 'sgd').run(scope, CUDAPlace(0)) — the "
                   "registered OpKernel wrappers; pageable numpy feeds, synchronous feed/run/fetch"}


def run_single(args):
    import torch
    from importlib import import_module
    pkg = load_pkg()
    capi, synth = pkg.capi, pkg.synth
    stepbench = import_module("paddle1_b200.stepbench")
    capi.lib()  # raises if the CUDA library is missing: no fallback
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (use --impl reference for the CPU arm)")
    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    v, d, n, a, opt = synth.CONFIGS["C2_ctr_sgd"]
    peak, peak_src = measured_peak_gbs()

    table_np = synth.table_values(v, d)
    table = torch.from_numpy(table_np).to(dev)
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
    bytes_ = synth.algorithmic_bytes(n, d, u_mean, opt)

    pipe = stepbench.SingleGpuPipeline(table, ids, dout, opt, 0.01)
    status, lr = pipe.status, pipe.lr
    pipe.capture()
    kernels_per_step = pipe.kernels_per_step()
    with torch.cuda.stream(pipe.main):
        pipe.run_steps(args.warmup)
        torch.cuda.synchronize()
        sampler = ClockSampler(0)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(pipe.main)
        pipe.run_steps(args.steps, args.warmup)
        e1.record(pipe.main)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        # keep the GPU under the same load a little longer so the clock sampler sees it
        t_end = time.perf_counter() + 0.25
        while time.perf_counter() < t_end:
            pipe.run_steps(50)
            torch.cuda.synchronize()
        clocks = sampler.stop()
    ms_per_step = ms / args.steps
    value = n * args.steps / (ms * 1e-3)

    # ---- per-phase timings (CUDA events on the launching stream, same rotating buffers) ----------
    reps = max(50, min(args.steps, 400))
    ph = pipe.phase_times(reps)
    t_gather, t_plan, t_update = ph["gather"], ph["sort_plan"], ph["update"]
    t_det = pipe.time_phase(lambda b: (pipe.gather(b), pipe.update(b, capi.DETERMINISTIC)), max(5, reps // 20))
    phases = {
        "gather_us": t_gather * 1e6, "sort_plan_us": t_plan * 1e6, "reduce_update_us": t_update * 1e6,
        "serial_step_us": ph["serial"] * 1e6,
        "gather_GBs": bytes_["fwd"] / t_gather / 1e9, "reduce_update_GBs": bytes_["upd"] / t_update / 1e9,
        "note": "each phase replayed alone as CUDA graphs over the rotating buffers; sort_plan moves 8 B per id per "
                "radix pass (not in the algorithmic bytes: SURVEY 8d counts it as overhead); serial_step = gather + "
                "sort + update back to back on one stream (what one OpKernel call sequence costs without the prefetch)",
    }
    cand = [("k_gather (lookup_table fwd)", bytes_["fwd"], t_gather),
            ("k_chunk_reduce (fused MergeAdd+SGD)", bytes_["upd"], t_update),
            ("sort plan (k_prepare_direct + k_radix_pass_direct x passes)", 0, t_plan)]
    hbm = [c for c in cand if c[1] > 0]
    dom = max(hbm, key=lambda c: c[2])
    achieved = dom[1] / dom[2] / 1e9
    # DRAM traffic of the same kernel(s) from the committed `ncu --set full` capture (profiles/), per launch
    traffic = None
    for fname in ("r2_ncu_full_summary.json", "r1_ncu_full_summary.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fname)) as f:
                prof = json.load(f)
            keys = ["k_gather"] if dom[0].startswith("k_gather") else ["k_chunk_reduce"]
            traffic = sum(v_["dram_traffic_bytes_per_launch"] for k, v_ in prof.items() if any(k.startswith(q) for q in keys))
            break
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": dom[0], "algorithmic_bytes_per_launch": dom[1],
                "peak_source": peak_src,
                "longest_phase": max(cand, key=lambda c: c[2])[0],
                "phases_frac": {"gather": bytes_["fwd"] / t_gather / 1e9 / peak,
                                "reduce_update": bytes_["upd"] / t_update / 1e9 / peak,
                                "serial_step (gather + sort plan + update, no overlap)": bytes_["step"] / ph["serial"] / 1e9 / peak},
                "step": {"algorithmic_bytes": bytes_["step"], "achieved": bytes_["step"] / (ms_per_step * 1e-3) / 1e9,
                         "frac": bytes_["step"] / (ms_per_step * 1e-3) / 1e9 / peak}}

    # ---- end to end: host buffers, H2D of ids+dOut and D2H of Out inside the timed region ---------
    e2e = run_e2e(torch, capi, table, lr, ids_h, dout_h, n, d, v, status, max(20, min(args.steps, 200)))
    try:
        e2e_ok = e2e_opkernel(torch, pkg, table_np, ids_h, dout_h, n, d, v, max(10, min(args.steps, 50)), True)
        e2e_ok["materialised_mode"] = e2e_opkernel(torch, pkg, table_np, ids_h, dout_h, n, d, v,
                                                   max(10, min(args.steps, 50)), False)
    except Exception as e:  # the wrappers are a separate library: report, do not hide
        e2e_ok = {"error": str(e)[:300]}

    st = int(status.item())
    parity = parity_c2(torch, capi, synth, dev)
    del pipe, table, ids, dout
    torch.cuda.empty_cache()
    extra = []
    if not getattr(args, "quick", False):
        free, _ = torch.cuda.mem_get_info()
        for name, need in (("C3_adam", 30e9), ("C5_dup_heavy", 100e9)):
            if free < need:
                extra.append({"config": name, "skipped": "needs %.0f GB of free device memory" % (need / 1e9)})
                continue
            extra.append(extra_config(torch, capi, synth, stepbench, name, dev, peak, 30 if name == "C3_adam" else 8))
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
        "roofline": roofline, "phases": phases, "det_ms_per_step": t_det * 1e3,
        "cpu_baseline": {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": e2e, "e2e_opkernel": e2e_ok, "parity": parity, "configs_extra": extra,
        "gpu_launches": kernels_per_step * args.steps, "clocks": clocks, "status_word": st,
    }
    print(json.dumps(line))
    if not parity["ok"] or any(not x.get("parity", {}).get("ok", True) for x in extra):
        raise SystemExit("parity check FAILED: %s" % json.dumps({"c2": parity, "extra": [x.get("parity") for x in extra]}))


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
                    help="sharded runs: ids per GPU per step (default 4096 x 26, the batch of configs[1])")
    ap.add_argument("--quick", action="store_true", help="skip configs_extra (C3 / C5) and the sharded sweep")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference_arm(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        if args.quick:
            os.environ["B200EMB_BENCH_QUICK"] = "1"
        import bench_sharded
        bench_sharded.run(args)
        return
    run_single(args)


if __name__ == "__main__":
    main()
