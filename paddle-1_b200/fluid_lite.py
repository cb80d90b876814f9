"""Python mirror of the handful of `paddle.fluid.core` / `paddle.fluid.op` names the reference's op unit
tests use (core.Scope, scope.var(name).get_tensor() / get_selected_rows(), tensor.set(np, place),
core.CPUPlace / core.CUDAPlace, Operator(type, **slots_and_attrs).run(scope, place)), bound by ctypes to
libb200emb_fluid.so: the OpKernel wrappers of paddle-1_b200/fluid/ registered with REGISTER_OP_CUDA_KERNEL
and dispatched like OperatorWithKernel::RunImpl does.  Lets tests/test_fluid_ops.py restate the reference's
own tests (python/paddle/fluid/tests/unittests/test_{lookup_table,sgd,adam,sum}_op.py) almost verbatim.

Only CUDA kernels exist: running an op on CPUPlace raises (no CPU fallback)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_CODES = {np.dtype(np.float32): 0, np.dtype(np.float64): 1, np.dtype(np.int64): 2, np.dtype(np.int32): 3}
_DTYPES = {v: k for k, v in _CODES.items()}

# slot names of the ops on the path (from the reference's OpMakers)
OP_SLOTS = {
    "lookup_table": (["W", "Ids"], ["Out"]),
    "lookup_table_grad": (["W", "Ids", "Out@GRAD"], ["W@GRAD"]),
    "sum": (["X"], ["Out"]),
    "sgd": (["Param", "Grad", "LearningRate"], ["ParamOut"]),
    "adam": (["Param", "Grad", "LearningRate", "Moment1", "Moment2", "Beta1Pow", "Beta2Pow"],
             ["ParamOut", "Moment1Out", "Moment2Out"]),
    "adagrad": (["Param", "Grad", "Moment", "LearningRate"], ["ParamOut", "MomentOut"]),
    "split_ids": (["Ids"], ["Out"]),
    "merge_ids": (["Ids", "X"], ["Out"]),
    "split_selected_rows": (["X"], ["Out"]),
    "extract_rows": (["X"], ["Out"]),
}
# attributes with their defaults (OpAttrChecker defaults: lookup_table_op.cc:71-83, adam_op.cc:99-112)
OP_ATTRS = {
    "lookup_table": {"is_sparse": ("b", False), "is_distributed": ("b", False), "padding_idx": ("l", -1)},
    "lookup_table_grad": {"is_sparse": ("b", False), "is_distributed": ("b", False), "padding_idx": ("l", -1)},
    "sum": {},
    "sgd": {},
    "adam": {"beta1": ("f", 0.9), "beta2": ("f", 0.999), "epsilon": ("f", 1.0e-8)},
    "adagrad": {"epsilon": ("f", 1.0e-6)},
    "split_ids": {}, "merge_ids": {}, "extract_rows": {},
    "split_selected_rows": {"height_sections": ("v", [])},
}


class FluidLiteError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libb200emb_fluid.so")
        if not os.path.exists(path):
            raise FluidLiteError("libb200emb_fluid.so not built (run __graft_entry__.build())")
        ctypes.CDLL(os.path.join(_HERE, "libb200emb.so"), mode=ctypes.RTLD_GLOBAL)
        _LIB = ctypes.CDLL(path)
        _LIB.fl_last_error.restype = ctypes.c_char_p
        _LIB.fl_session_new.restype = ctypes.c_void_p
        _LIB.fl_pe_new.restype = ctypes.c_void_p
        _LIB.fl_pe_scope.restype = ctypes.c_void_p
        _LIB.fl_sr_height.restype = ctypes.c_int64
        _LIB.fl_sr_num_rows.restype = ctypes.c_int64
    return _LIB


def _check(rc):
    if rc < 0:
        raise FluidLiteError(lib().fl_last_error().decode())
    return rc


class CPUPlace:
    device = -1


class CUDAPlace:
    def __init__(self, device=0):
        self.device = int(device)


def is_compiled_with_cuda():
    return True


class Tensor:
    def __init__(self, scope, name):
        self._s, self._n = scope, name.encode()

    def set(self, array, place):
        a = np.ascontiguousarray(array)
        if a.dtype not in _CODES:
            raise TypeError("unsupported dtype %s" % a.dtype)
        dims = (ctypes.c_int64 * max(a.ndim, 1))(*a.shape)
        _check(lib().fl_tensor_set(self._s._h, self._n, _CODES[a.dtype], dims, a.ndim,
                                   a.ctypes.data_as(ctypes.c_void_p), place.device))

    def shape(self):
        dims = (ctypes.c_int64 * 16)()
        code = ctypes.c_int(0)
        nd = lib().fl_tensor_info(self._s._h, self._n, dims, ctypes.byref(code))
        if nd == -1:
            _check(-1)
        return None if nd < 0 else (list(dims[:nd]), code.value)

    def __array__(self, dtype=None, copy=None):
        info = self.shape()
        if info is None:
            raise FluidLiteError("tensor %s holds no data" % self._n.decode())
        dims, code = info
        out = np.empty(dims, _DTYPES[code])
        _check(lib().fl_tensor_get(self._s._h, self._n, out.ctypes.data_as(ctypes.c_void_p)))
        return out if dtype is None else out.astype(dtype)


class SelectedRows:
    def __init__(self, scope, name):
        self._s, self._n = scope, name.encode()

    def set_rows(self, rows):
        r = np.ascontiguousarray(rows, dtype=np.int64)
        _check(lib().fl_sr_set_rows(self._s._h, self._n, r.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(r.size)))

    def set_height(self, h):
        _check(lib().fl_sr_set_height(self._s._h, self._n, ctypes.c_int64(h)))

    def height(self):
        return int(lib().fl_sr_height(self._s._h, self._n))

    def rows(self):
        n = int(_check(lib().fl_sr_num_rows(self._s._h, self._n)))
        out = np.empty(n, np.int64)
        _check(lib().fl_sr_get_rows(self._s._h, self._n, out.ctypes.data_as(ctypes.c_void_p)))
        return out.tolist()

    def get_tensor(self):
        return Tensor(self._s, self._n.decode())


class Variable:
    def __init__(self, scope, name):
        self._s, self._name = scope, name

    def get_tensor(self):
        if lib().fl_var_kind(self._s._h, self._name.encode()) == 0:
            _check(lib().fl_var_tensor(self._s._h, self._name.encode()))
        return Tensor(self._s, self._name)

    def get_selected_rows(self):
        _check(lib().fl_var_selected_rows(self._s._h, self._name.encode()))
        return SelectedRows(self._s, self._name)


class Scope:
    def __init__(self, handle=None):
        self._owned = handle is None
        self._h = ctypes.c_void_p(lib().fl_session_new() if handle is None else handle)

    def __del__(self):
        try:
            if self._owned:
                lib().fl_session_free(self._h)
        except Exception:
            pass

    def var(self, name):
        return Variable(self, name)

    def find_var(self, name):
        return Variable(self, name)


class Operator:
    """Operator("sgd", Param='Param', Grad='Grad', ParamOut='Param', LearningRate='LearningRate')"""

    def __init__(self, op_type, **kwargs):
        if op_type not in OP_SLOTS:
            raise FluidLiteError("operator %s is not part of the sparse-embedding path" % op_type)
        self.type = op_type
        ins, outs = OP_SLOTS[op_type]
        self.inputs, self.outputs, self.attrs = {}, {}, {}
        for k, v in kwargs.items():
            names = [v] if isinstance(v, str) else list(v) if isinstance(v, (list, tuple)) else None
            if k in ins:
                self.inputs[k] = names
            elif k in outs:
                self.outputs[k] = names
            elif k in OP_ATTRS[op_type]:
                self.attrs[k] = v
            else:
                raise FluidLiteError("%s is neither a slot nor an attribute of %s" % (k, op_type))

    def run(self, scope, place):
        def slots(d):
            return ";".join("%s=%s" % (k, ",".join(v)) for k, v in d.items()).encode()
        attrs = []
        for name, (kind, default) in OP_ATTRS[self.type].items():
            val = self.attrs.get(name, default)
            if kind == "b":
                val = "1" if val else "0"
            elif kind == "f":
                val = repr(float(val))
            elif kind == "v":
                val = "|".join(str(int(x)) for x in val)
            attrs.append("%s:%s:%s" % (name, kind, val))
        _check(lib().fl_op_run(scope._h, self.type.encode(), slots(self.inputs), slots(self.outputs),
                               ";".join(attrs).encode(), place.device))


def set_fused(on):
    """FLAGS_b200emb_fused: the sparse gradient stays on the device between lookup_table_grad, sum and the
    optimizer op (no rows() round trip, no copies) and lookup_table starts the sort plan early."""
    lib().fl_set_fused(1 if on else 0)


def registered_kernels(op_type):
    return int(lib().fl_num_registered_kernels(op_type.encode()))


class ShardedParallelSession:
    """The slice of ParallelExecutor the sharded-table op handles need (fluid/sparse_all_to_all_op_handle.h):
    one local scope per device (a device id may repeat: several ranks on one GPU, like the fake places of the
    reference's op-handle tests), the pass's set-up for one embedding parameter, and one Run of each handle."""

    SGD, ADAM = 0, 1

    def __init__(self, devices):
        arr = (ctypes.c_int * len(devices))(*devices)
        self._h = ctypes.c_void_p(lib().fl_pe_new(len(devices), arr))
        self.devices = list(devices)
        self.scopes = [Scope(lib().fl_pe_scope(self._h, i)) for i in range(len(devices))]

    def __del__(self):
        try:
            lib().fl_pe_free(self._h)
        except Exception:
            pass

    def shard_table(self, param, height, d, max_ids, max_recv, optimizer="sgd", moment1="", moment2=""):
        _check(lib().fl_pe_shard_table(self._h, param.encode(), moment1.encode(), moment2.encode(),
                                       ctypes.c_int64(height), ctypes.c_int64(d),
                                       self.ADAM if optimizer == "adam" else self.SGD,
                                       ctypes.c_int64(max_ids), ctypes.c_int64(max_recv)))

    def sharded_lookup(self, ids, out):
        _check(lib().fl_pe_sharded_lookup(self._h, ids.encode(), out.encode()))

    def sparse_all_to_all(self, grad, lr, beta1_pow="", beta2_pow="", beta1=0.9, beta2=0.999, epsilon=1e-8):
        _check(lib().fl_pe_sparse_all_to_all(self._h, grad.encode(), lr.encode(), beta1_pow.encode(), beta2_pow.encode(),
                                             ctypes.c_float(beta1), ctypes.c_float(beta2), ctypes.c_float(epsilon)))

    def sync(self):
        """Waits for every device; returns the OR of the engines' status words."""
        return _check(lib().fl_pe_sync(self._h))
