#!/usr/bin/env python
"""Generates tests/golden/ref_outputs.npz: outputs of the REFERENCE's own CPU kernels (oracle/_ref — compiled from
the unmodified sources under /root/reference by oracle/Makefile.ref) on seeded random inputs, so that the oracle
restatement stays pinned to the reference on machines that have neither /root/reference nor the built oracle/_ref.

    python tests/golden/make_ref_fixtures.py        (in the container that mounts /root/reference)

Inputs are NOT stored: `inputs(case)` below regenerates them from the case's seed; tests/test_oracle_goldens.py
imports this module for that.  Outputs are stored as raw bit patterns (uint32 / uint64 views)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "ref_outputs.npz")

# (seed, n, v, d): small enough for a fixture of a few hundred KB, shaped like the cases of test_ref_vs_oracle.py —
# a hot row (a third of the batch), duplicates, widths that are and are not multiples of 4
CASES = [(11, 1, 3, 4), (12, 50, 9, 7), (13, 300, 40, 33), (14, 600, 200, 16)]
DTYPES = [np.float32, np.float64]


def inputs(seed, n, v, d, dtype):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, v, size=n).astype(np.int64)
    rows[: n // 3] = rows[0]
    value = (rng.standard_normal((n, d)) * 0.01).astype(dtype)
    table = (rng.random((v, d)) * 0.1 - 0.05).astype(dtype)
    m1 = (rng.standard_normal((v, d)) * 1e-3).astype(dtype)
    m2 = (rng.random((v, d)) * 1e-4).astype(dtype)
    lr = np.array([0.01], dtype)
    b1p, b2p = np.array([0.9 ** 3], dtype), np.array([0.999 ** 3], dtype)
    return {"rows": rows, "value": value, "table": table, "m1": m1, "m2": m2, "lr": lr, "b1p": b1p, "b2p": b2p}


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64) if a.dtype.kind == "f" else a


def outputs(impl, x, adam_kwargs=None):
    """Every kernel of the path on one input set; `impl` is oracle.ref_cpu (generation) or oracle (the test)."""
    out = {}
    out["lookup"] = impl.lookup_table(x["table"], x["rows"])
    out["lookup_pad"] = impl.lookup_table(x["table"], x["rows"], int(x["rows"][-1]))
    r, v = impl.merge_add(x["rows"], x["value"])
    out["merge_rows"], out["merge_value"] = r, v
    t = x["table"].copy()
    impl.selected_rows_add_to_tensor(x["rows"], x["value"], t)
    out["add_to_tensor"] = t
    p = x["table"].copy()
    impl.sgd_sparse(p, x["rows"], x["value"], x["lr"])
    out["sgd"] = p
    a = [x["table"].copy(), x["m1"].copy(), x["m2"].copy()]
    impl.adam_sparse(a[0], a[1], a[2], x["rows"], x["value"], x["lr"], x["b1p"], x["b2p"], **(adam_kwargs or {}))
    out["adam_p"], out["adam_m1"], out["adam_m2"] = a
    return out


def key(case, dtype, name):
    return "s%d_n%d_v%d_d%d_%s_%s" % (*case, np.dtype(dtype).name, name)


def main():
    sys.path.insert(0, ROOT)
    from oracle import ref_cpu
    if not ref_cpu.available():
        raise SystemExit("oracle/_ref is not built: run oracle.build(ref=True) where /root/reference is mounted")
    store = {}
    for case in CASES:
        for dt in DTYPES:
            x = inputs(*case, dt)
            for name, arr in outputs(ref_cpu, x).items():
                store[key(case, dt, name)] = bits(arr)
    np.savez_compressed(OUT, **store)
    print("wrote %s: %d arrays, %d bytes" % (OUT, len(store), os.path.getsize(OUT)))


if __name__ == "__main__":
    main()
