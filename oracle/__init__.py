"""numpy/ctypes front-end of the CPU oracle (oracle/oracle.cc).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package never
imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_I64P = ctypes.POINTER(ctypes.c_int64)
_VP = ctypes.c_void_p


def build(ref=False):
    """Compile liboracle.so (and oracle/_ref when ref=True and the reference tree is mounted)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "all"])
    if ref and os.path.isdir("/root/reference/paddle/fluid/operators"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_abi_version.restype = ctypes.c_int
    return _LIB


def _sfx(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return "f32"
    if dtype == np.float64:
        return "f64"
    raise TypeError("oracle supports float32/float64, got %s" % dtype)


def _p(a):
    return a.ctypes.data_as(_VP)


def _ids(a):
    a = np.ascontiguousarray(a, dtype=np.int64).reshape(-1)
    return a


def _fn(name, restype=ctypes.c_int64):
    f = getattr(lib(), name)
    f.restype = restype
    return f


def lookup_table(table, ids, padding_idx=-1):
    """Out[k,:] = W[ids[k],:], padding rows zero (lookup_table_op.h:36-84)."""
    table = np.ascontiguousarray(table)
    ids = _ids(ids)
    v, d = table.shape
    out = np.empty((ids.size, d), dtype=table.dtype)
    rc = _fn("oracle_lookup_table_" + _sfx(table.dtype))(
        _p(table), ctypes.c_int64(v), ctypes.c_int64(d), _p(ids), ctypes.c_int64(ids.size),
        ctypes.c_int64(padding_idx), _p(out))
    if rc != 0:
        raise IndexError("lookup_table: ids[%d]=%d outside [0,%d)" % (-rc - 1, ids[-rc - 1], v))
    return out


def lookup_table_selected_rows(table_rows, table_value, ids, padding_idx=-1):
    table_value = np.ascontiguousarray(table_value)
    table_rows = _ids(table_rows)
    ids = _ids(ids)
    d = table_value.shape[1]
    out = np.empty((ids.size, d), dtype=table_value.dtype)
    rc = _fn("oracle_lookup_table_selected_rows_" + _sfx(table_value.dtype))(
        _p(table_rows), ctypes.c_int64(table_rows.size), _p(table_value), ctypes.c_int64(d),
        _p(ids), ctypes.c_int64(ids.size), ctypes.c_int64(padding_idx), _p(out))
    if rc != 0:
        raise IndexError("lookup_table(SelectedRows W): ids[%d] missing" % (-rc - 1))
    return out


def lookup_table_grad_sparse(ids, d_out):
    """-> (rows, value): rows = ids verbatim, value = copy of dOut (lookup_table_op.h:106-134)."""
    d_out = np.ascontiguousarray(d_out)
    ids = _ids(ids)
    d = d_out.shape[-1]
    rows = np.empty(ids.size, dtype=np.int64)
    value = np.empty((ids.size, d), dtype=d_out.dtype)
    _fn("oracle_lookup_table_grad_sparse_" + _sfx(d_out.dtype), None)(
        _p(ids), ctypes.c_int64(ids.size), _p(d_out), ctypes.c_int64(d), _p(rows), _p(value))
    return rows, value


def lookup_table_grad_dense(ids, d_out, height):
    d_out = np.ascontiguousarray(d_out)
    ids = _ids(ids)
    d = d_out.shape[-1]
    dt = np.empty((height, d), dtype=d_out.dtype)
    rc = _fn("oracle_lookup_table_grad_dense_" + _sfx(d_out.dtype))(
        _p(ids), ctypes.c_int64(ids.size), _p(d_out), ctypes.c_int64(height), ctypes.c_int64(d), _p(dt))
    if rc != 0:
        raise IndexError("lookup_table_grad: bad id at %d" % (-rc - 1))
    return dt


def sum_selected_rows(inputs):
    """`sum` op, SelectedRows output (sum_op.h:71-145): concat rows/values, skip empties.

    inputs: list of (rows, value[n_i, d]).  Returns (rows, value)."""
    total = sum(len(_ids(r)) for r, _ in inputs)
    first = next((v for r, v in inputs if len(_ids(r)) > 0), inputs[-1][1])
    first = np.asarray(first)
    d = first.shape[1] if first.ndim == 2 else 0
    out_rows = np.empty(total, dtype=np.int64)
    out_value = np.empty((total, d), dtype=first.dtype)
    if total == 0:
        return out_rows, out_value
    fill = 0
    offset = 0
    for r, v in inputs:
        r = _ids(r)
        if r.size == 0:
            continue
        v = np.ascontiguousarray(v, dtype=first.dtype)
        flat = out_value.reshape(-1)
        _fn("oracle_selected_rows_add_to_" + _sfx(first.dtype), None)(
            _p(r), ctypes.c_int64(r.size), _p(v), ctypes.c_int64(d), _p(out_rows),
            ctypes.c_int64(fill), _p(flat), ctypes.c_int64(offset))
        fill += r.size
        offset += v.size
    return out_rows, out_value


def selected_rows_add_to_tensor(rows, value, tensor):
    """tensor[rows[i],:] += value[i,:] in input order, in place (selected_rows_functor.cc:153-178)."""
    rows = _ids(rows)
    value = np.ascontiguousarray(value)
    assert tensor.flags["C_CONTIGUOUS"] and tensor.dtype == value.dtype
    _fn("oracle_selected_rows_add_to_tensor_" + _sfx(value.dtype), None)(
        _p(rows), ctypes.c_int64(rows.size), _p(value), ctypes.c_int64(value.shape[1]), _p(tensor))
    return tensor


def selected_rows_add_tensor(rows, value, in2):
    rows = _ids(rows)
    value = np.ascontiguousarray(value)
    in2 = np.ascontiguousarray(in2, dtype=value.dtype)
    out = np.empty_like(in2)
    _fn("oracle_selected_rows_add_tensor_" + _sfx(value.dtype), None)(
        _p(rows), ctypes.c_int64(rows.size), _p(value), ctypes.c_int64(value.shape[1]),
        ctypes.c_int64(in2.shape[0]), _p(in2), _p(out))
    return out


def merge_add(rows, value, reference_form=False, truth=False):
    """scatter::MergeAdd (selected_rows_functor.cc:193-228) -> (merged_rows[U], merged_value[U,d]).

    reference_form=True uses the reference's O(N*U) linear FindPos; the default is
    bit-identical (same add order) but hash-indexed.  truth=True accumulates in fp64."""
    rows = _ids(rows)
    value = np.ascontiguousarray(value)
    d = value.shape[1]
    out_rows = np.empty(rows.size, dtype=np.int64)
    if truth:
        out_value = np.empty((rows.size, d), dtype=np.float64)
        name = "oracle_merge_add_truth_"
    else:
        out_value = np.empty((rows.size, d), dtype=value.dtype)
        name = "oracle_merge_add_ref_" if reference_form else "oracle_merge_add_"
    u = _fn(name + _sfx(value.dtype))(
        _p(rows), ctypes.c_int64(rows.size), _p(value), ctypes.c_int64(d), _p(out_rows), _p(out_value))
    return out_rows[:u].copy(), out_value[:u].copy()


def sgd_sparse(param, rows, grad, lr):
    """Sparse SGD in place, no merge, per-occurrence (sgd_op.h:47-81)."""
    rows = _ids(rows)
    grad = np.ascontiguousarray(grad, dtype=param.dtype)
    lr = np.ascontiguousarray(lr, dtype=param.dtype).reshape(-1)
    assert param.flags["C_CONTIGUOUS"]
    v, d = param.shape
    rc = _fn("oracle_sgd_sparse_" + _sfx(param.dtype))(
        _p(param), ctypes.c_int64(v), ctypes.c_int64(d), _p(rows), ctypes.c_int64(rows.size),
        _p(grad), _p(lr))
    if rc != 0:
        raise IndexError("sgd: rows[%d] >= height" % (-rc - 1))
    return param


def adam_sparse(param, mom1, mom2, rows, grad, lr, beta1_pow, beta2_pow,
                beta1=0.9, beta2=0.999, epsilon=1e-8, reference_form=False):
    """Sparse (lazy) Adam in place = MergeAdd + SparseAdamFunctor (adam_op.h:199-219,282-321)."""
    rows = _ids(rows)
    dt = param.dtype
    grad = np.ascontiguousarray(grad, dtype=dt)
    lr = np.ascontiguousarray(lr, dtype=dt).reshape(-1)
    b1p = np.ascontiguousarray(beta1_pow, dtype=dt).reshape(-1)
    b2p = np.ascontiguousarray(beta2_pow, dtype=dt).reshape(-1)
    for a in (param, mom1, mom2):
        assert a.flags["C_CONTIGUOUS"] and a.dtype == dt
    v, d = param.shape
    u = _fn("oracle_adam_sparse_" + _sfx(dt))(
        _p(param), _p(mom1), _p(mom2), ctypes.c_int64(v), ctypes.c_int64(d), _p(rows),
        ctypes.c_int64(rows.size), _p(grad), _p(lr), _p(b1p), _p(b2p),
        ctypes.c_float(beta1), ctypes.c_float(beta2), ctypes.c_float(epsilon),
        ctypes.c_int(1 if reference_form else 0))
    return u


def adagrad_sparse(param, moment, rows, grad, lr, epsilon):
    rows = _ids(rows)
    dt = param.dtype
    grad = np.ascontiguousarray(grad, dtype=dt)
    lr = np.ascontiguousarray(lr, dtype=dt).reshape(-1)
    v, d = param.shape
    return _fn("oracle_adagrad_sparse_" + _sfx(dt))(
        _p(param), _p(moment), ctypes.c_int64(v), ctypes.c_int64(d), _p(rows),
        ctypes.c_int64(rows.size), _p(grad), _p(lr), ctypes.c_float(epsilon))


def split_ids(ids, shard_num):
    """-> list of per-shard id arrays, shard = id % shard_num, input order (split_ids_op.h:36-58)."""
    ids = _ids(ids)
    out = np.empty(ids.size, dtype=np.int64)
    counts = np.zeros(shard_num, dtype=np.int64)
    _fn("oracle_split_ids", None)(_p(ids), ctypes.c_int64(ids.size), ctypes.c_int64(shard_num),
                                  _p(out), _p(counts))
    offs = np.concatenate([[0], np.cumsum(counts)])
    return [out[offs[s]:offs[s + 1]].copy() for s in range(shard_num)]


def merge_ids(ids, shard_rows):
    """Inverse of split_ids on row blocks (merge_ids_op.h:63-87). shard_rows: list of [n_s, d] float32."""
    ids = _ids(ids)
    shard_num = len(shard_rows)
    d = max(int(x.shape[1]) for x in shard_rows if x.ndim == 2)
    blocks = [np.ascontiguousarray(x, dtype=np.float32).reshape(-1, d) for x in shard_rows]
    x = np.concatenate(blocks, axis=0) if blocks else np.empty((0, d), np.float32)
    offs = np.zeros(shard_num, dtype=np.int64)
    offs[1:] = np.cumsum([b.shape[0] for b in blocks])[:-1]
    out = np.empty((ids.size, d), dtype=np.float32)
    _fn("oracle_merge_ids_f32", None)(_p(ids), ctypes.c_int64(ids.size), ctypes.c_int64(shard_num),
                                      _p(x), _p(offs), ctypes.c_int64(d), _p(out))
    return out


def split_ids_selected_rows(rows, value, shard_num):
    rows = _ids(rows)
    value = np.ascontiguousarray(value, dtype=np.float32)
    d = value.shape[1]
    out_rows = np.empty(rows.size, dtype=np.int64)
    out_value = np.empty((rows.size, d), dtype=np.float32)
    counts = np.zeros(shard_num, dtype=np.int64)
    _fn("oracle_split_ids_selected_rows_f32", None)(
        _p(rows), ctypes.c_int64(rows.size), _p(value), ctypes.c_int64(d),
        ctypes.c_int64(shard_num), _p(out_rows), _p(out_value), _p(counts))
    offs = np.concatenate([[0], np.cumsum(counts)])
    return [(out_rows[offs[s]:offs[s + 1]].copy(), out_value[offs[s]:offs[s + 1]].copy())
            for s in range(shard_num)]


def split_selected_rows(rows, value, height_sections):
    rows = _ids(rows)
    value = np.ascontiguousarray(value, dtype=np.float32)
    d = value.shape[1]
    hs = np.ascontiguousarray(height_sections, dtype=np.int64)
    out_rows = np.empty(rows.size, dtype=np.int64)
    out_value = np.empty((rows.size, d), dtype=np.float32)
    counts = np.zeros(hs.size, dtype=np.int64)
    _fn("oracle_split_selected_rows_f32", None)(
        _p(rows), ctypes.c_int64(rows.size), _p(value), ctypes.c_int64(d), _p(hs),
        ctypes.c_int64(hs.size), _p(out_rows), _p(out_value), _p(counts))
    offs = np.concatenate([[0], np.cumsum(counts)])
    return [(out_rows[offs[s]:offs[s + 1]].copy(), out_value[offs[s]:offs[s + 1]].copy())
            for s in range(hs.size)]


def extract_rows(rows):
    """extract_rows_op.cc:49-77: Out[n,1] = the SelectedRows' row ids (duplicates kept)."""
    return _ids(rows).reshape(-1, 1).copy()


def sparse_decay(table, rows, coeff, mode=0, addend=None):
    """The regularizers' SelectedRows op chain (regularizer.py:146-170 / 226-253 + 64-69):
    out = [addend +] coeff * f(table[rows]), f = identity (mode 0, L2) or sign (mode 1, L1)."""
    table = np.ascontiguousarray(table)
    rows = _ids(rows)
    v, d = table.shape
    out = np.empty((rows.size, d), dtype=table.dtype)
    ct = ctypes.c_float if table.dtype == np.float32 else ctypes.c_double
    add = None if addend is None else np.ascontiguousarray(addend, dtype=table.dtype)
    rc = _fn("oracle_sparse_decay_" + _sfx(table.dtype))(
        _p(table), ctypes.c_int64(v), ctypes.c_int64(d), _p(rows), ctypes.c_int64(rows.size), ct(coeff),
        ctypes.c_int(mode), None if add is None else _p(add), _p(out))
    if rc != 0:
        raise IndexError("sparse_decay: rows[%d] outside [0,%d)" % (-rc - 1, v))
    return out


def gather_selected_rows(per_device):
    """ReduceOpHandle's SelectedRows branch = concat in device order (reduce_and_gather.h:56-93)."""
    return sum_selected_rows(per_device)
