"""Builds libb200emb_fluid.so: the OpKernel wrappers + the mini op runner, compiled for sm_100a against
the repo's compat/ headers and linked to libb200emb.so (same directory, $ORIGIN rpath)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
OUT = os.path.join(PKG, "libb200emb_fluid.so")
SOURCES = ["lookup_table_op.cu", "sum_op.cu", "sgd_op.cu", "adam_op.cu", "adagrad_op.cu", "split_ids_op.cu",
           "split_selected_rows_op.cu", "sparse_all_to_all_op_handle.cu", "runner.cu"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    srcs = [os.path.join(HERE, s) for s in SOURCES]
    deps = list(srcs) + [os.path.join(HERE, h) for h in ("b200_op_common.h", "b200_side_table.h", "sparse_all_to_all_op_handle.h")] + [os.path.join(PKG, "libb200emb.so")]
    for base, _, files in os.walk(os.path.join(PKG, "compat")):
        deps += [os.path.join(base, f) for f in files]
    if not (force or _newer(OUT, deps)):
        return OUT
    nvcc = os.environ.get("NVCC") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc")
    host_cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
           "-ccbin", host_cxx, "-DPADDLE_WITH_CUDA", "-Xcompiler", "-fPIC", "-shared",
           "-I", os.path.join(PKG, "compat"), "-I", os.path.join(ROOT, "include"), "-I", HERE,
           "-o", OUT] + srcs + ["-L", PKG, "-lb200emb", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN"]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    build(force=True, verbose=True)
