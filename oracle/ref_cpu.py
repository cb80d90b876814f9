"""numpy front-end of oracle/_ref/libref_cpu.so — the reference's OWN CPU kernels for this path,
compiled unmodified from /root/reference by `make -C oracle ref` (see oracle/ref_driver.cc).

TEST / BASELINE INFRASTRUCTURE ONLY.  Used to validate the restated oracle bit for bit and, when
present, as bench.py's CPU arm (cpu_baseline.kind = "reference").  Same call signatures as the
functions of the same name in oracle/__init__.py."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libref_cpu.so")
_LIB = None
_VP = ctypes.c_void_p
_I64 = ctypes.c_int64


def available():
    return os.path.exists(_PATH)


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(_PATH)
        _LIB.ref_last_error.restype = ctypes.c_char_p
    return _LIB


def _sfx(dt):
    return "f32" if np.dtype(dt) == np.float32 else "f64"


def _p(a):
    return a.ctypes.data_as(_VP)


def _ids(a):
    return np.ascontiguousarray(a, dtype=np.int64).reshape(-1)


def _fn(name, restype=ctypes.c_int):
    f = getattr(lib(), name)
    f.restype = restype
    return f


def _err(what):
    raise RuntimeError("%s: %s" % (what, lib().ref_last_error().decode()))


def lookup_table(table, ids, padding_idx=-1):
    table = np.ascontiguousarray(table)
    ids = _ids(ids)
    v, d = table.shape
    out = np.empty((ids.size, d), table.dtype)
    if _fn("ref_lookup_table_" + _sfx(table.dtype))(_p(table), _I64(v), _I64(d), _p(ids), _I64(ids.size),
                                                   _I64(padding_idx), _p(out)) != 0:
        raise IndexError(lib().ref_last_error().decode())
    return out


def lookup_table_grad_sparse(ids, d_out, height=None):
    d_out = np.ascontiguousarray(d_out)
    ids = _ids(ids)
    d = d_out.shape[-1]
    height = int(height if height is not None else (ids.max() + 1 if ids.size else 1))
    rows = np.empty(ids.size, np.int64)
    value = np.empty((ids.size, d), d_out.dtype)
    if _fn("ref_lookup_table_grad_sparse_" + _sfx(d_out.dtype))(
            _p(ids), _I64(ids.size), _p(d_out), _I64(d), _I64(height), _p(rows), _p(value)) != 0:
        _err("lookup_table_grad")
    return rows, value


def sum_selected_rows(inputs, height=1 << 40):
    k = len(inputs)
    rows = [_ids(r) for r, _ in inputs]
    first = next((np.asarray(v) for r, v in zip(rows, (v for _, v in inputs)) if r.size), np.asarray(inputs[-1][1]))
    d = first.shape[1]
    vals = [np.ascontiguousarray(v, dtype=first.dtype).reshape(-1, d) for _, v in inputs]
    total = sum(r.size for r in rows)
    out_rows = np.empty(total, np.int64)
    out_value = np.empty((total, d), first.dtype)
    rp = (_VP * k)(*[_p(r) for r in rows])
    vp = (_VP * k)(*[_p(v) for v in vals])
    ns = np.array([r.size for r in rows], np.int64)
    got = _fn("ref_sum_selected_rows_" + _sfx(first.dtype), ctypes.c_int64)(
        ctypes.c_int(k), rp, _p(ns), vp, _I64(d), _I64(height), _p(out_rows), _p(out_value))
    if got < 0:
        _err("sum")
    return out_rows[:got], out_value[:got]


def selected_rows_add_to_tensor(rows, value, tensor):
    rows = _ids(rows)
    value = np.ascontiguousarray(value)
    if _fn("ref_selected_rows_add_to_tensor_" + _sfx(value.dtype))(
            _p(rows), _I64(rows.size), _p(value), _I64(value.shape[1]), _I64(tensor.shape[0]), _p(tensor)) != 0:
        _err("SelectedRowsAddToTensor")
    return tensor


def merge_add(rows, value, height=1 << 40):
    rows = _ids(rows)
    value = np.ascontiguousarray(value)
    d = value.shape[1]
    out_rows = np.empty(rows.size, np.int64)
    out_value = np.empty((rows.size, d), value.dtype)
    u = _fn("ref_merge_add_" + _sfx(value.dtype), ctypes.c_int64)(
        _p(rows), _I64(rows.size), _p(value), _I64(d), _I64(height), _p(out_rows), _p(out_value))
    if u < 0:
        _err("MergeAdd")
    return out_rows[:u].copy(), out_value[:u].copy()


def sgd_sparse(param, rows, grad, lr):
    rows = _ids(rows)
    grad = np.ascontiguousarray(grad, dtype=param.dtype)
    lr = np.ascontiguousarray(lr, dtype=param.dtype).reshape(-1)
    v, d = param.shape
    if _fn("ref_sgd_sparse_" + _sfx(param.dtype))(_p(param), _I64(v), _I64(d), _p(rows), _I64(rows.size),
                                                 _p(grad), _p(lr)) != 0:
        raise IndexError(lib().ref_last_error().decode())
    return param


def adam_sparse(param, mom1, mom2, rows, grad, lr, beta1_pow, beta2_pow, beta1=0.9, beta2=0.999, epsilon=1e-8):
    rows = _ids(rows)
    dt = param.dtype
    grad = np.ascontiguousarray(grad, dtype=dt)
    lr = np.ascontiguousarray(lr, dtype=dt).reshape(-1)
    b1p = np.ascontiguousarray(beta1_pow, dtype=dt).reshape(-1)
    b2p = np.ascontiguousarray(beta2_pow, dtype=dt).reshape(-1)
    v, d = param.shape
    if _fn("ref_adam_sparse_" + _sfx(dt))(_p(param), _p(mom1), _p(mom2), _I64(v), _I64(d), _p(rows),
                                         _I64(rows.size), _p(grad), _p(lr), _p(b1p), _p(b2p),
                                         ctypes.c_float(beta1), ctypes.c_float(beta2), ctypes.c_float(epsilon)) != 0:
        _err("adam")
