"""paddle-1_b200: B200 (sm_100a) kernels for PaddlePaddle Fluid's sparse-embedding hot path.

Layout
  csrc/      hand-written CUDA kernels + the C ABI (include/b200emb.h) -> libb200emb.so
  compat/    own headers mirroring the Fluid API subset the op kernels touch
  fluid/     OpKernel wrappers registered with REGISTER_OP_CUDA_KERNEL + a mini op runner
  capi.py    ctypes binding of the C ABI over torch CUDA tensors (memory + streams only)
  synth.py   synthetic id streams (bounded Zipf) of BASELINE.json's configs

There is no CPU fallback: every op raises if libb200emb.so is missing or no CUDA device exists.
"""
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "libb200emb.so")
FLUID_LIB_PATH = os.path.join(PKG_DIR, "libb200emb_fluid.so")

NVCC_FLAGS = ["-O3", "-std=c++20", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-fmad=false", "-Xcompiler", "-fPIC", "-shared"]
HEADERS = ["common.cuh", "internal.h", "segreduce.cuh"]
KERNEL_SOURCES = ["capi.cu", "sort.cu", "cluster_sort.cu", "gather.cu", "segreduce_f32.cu", "segreduce_f64.cu", "shard.cu", "regularize.cu", "peer.cu", "noalloc.cu", "shard_engine.cu"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _compile_objects(nvcc, flags, srcs, objdir, deps, force, verbose):
    """One nvcc -c per translation unit, in parallel; an object is rebuilt when its source or a header is newer."""
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(objdir, exist_ok=True)
    jobs, objs = [], []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if force or _stale(o, [s] + deps):
            jobs.append([nvcc] + flags + ["-c", s, "-o", o])
    if verbose:
        for j in jobs:
            print(" ".join(j), file=sys.stderr)
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        list(ex.map(subprocess.check_call, jobs))
    return objs, bool(jobs)


def build_native(force=False, verbose=False):
    """Compile libb200emb.so (and libb200emb_fluid.so when its sources exist) in-tree for sm_100a."""
    csrc = os.path.join(PKG_DIR, "csrc")
    srcs = [os.path.join(csrc, s) for s in KERNEL_SOURCES]
    headers = [os.path.join(csrc, h) for h in HEADERS] + [os.path.join(ROOT, "include", "b200emb.h")]
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]
    objs, rebuilt = _compile_objects(_nvcc(), compile_flags, srcs, os.path.join(PKG_DIR, "build", "csrc"),
                                     headers, force, verbose)
    if rebuilt or _stale(LIB_PATH, objs):
        cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH] + objs
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.check_call(cmd)
    fluid_dir = os.path.join(PKG_DIR, "fluid")
    build_script = os.path.join(fluid_dir, "build.py")
    if os.path.exists(build_script):
        import importlib.util
        spec = importlib.util.spec_from_file_location("paddle1_b200_fluid_build", build_script)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build(force=force, verbose=verbose)
    return LIB_PATH


from . import capi  # noqa: E402
from . import synth  # noqa: E402
from . import sharded  # noqa: E402
