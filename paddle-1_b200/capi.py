"""ctypes binding of include/b200emb.h over torch CUDA tensors.

torch is used for device memory and streams only; every computation is a kernel of
libb200emb.so.  Missing library or missing GPU raises — there is no fallback path.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libb200emb.so")

DETERMINISTIC = 1
PLAN_READY = 2
UNIQUE_READY = 4

_ERRORS = {-1: "invalid argument", -2: "workspace missing/misaligned/too small",
           -3: "CUDA error", -4: "size beyond supported limits",
           -5: "request needs more than the library's built-in 8 MiB scratch"}
NOALLOC_SCRATCH_BYTES = 8 * 1024 * 1024

EXPORTS = [
    "b200emb_abi_version", "b200emb_last_cuda_error", "b200emb_workspace_bytes",
    "b200emb_lookup_table_f32", "b200emb_lookup_table_f64",
    "b200emb_lookup_table_grad_sparse_f32", "b200emb_lookup_table_grad_sparse_f64",
    "b200emb_selected_rows_add_to", "b200emb_plan_rows",
    "b200emb_merge_add_f32", "b200emb_merge_add_f64",
    "b200emb_selected_rows_add_to_tensor_f32", "b200emb_selected_rows_add_to_tensor_f64",
    "b200emb_sparse_sgd_f32", "b200emb_sparse_sgd_f64",
    "b200emb_sparse_adam_f32", "b200emb_sparse_adam_f64",
    "b200emb_sparse_sgd_noalloc_f32", "b200emb_sparse_adagrad_f32", "b200emb_sparse_adagrad_f64",
    "b200emb_plan_unique_rows", "b200emb_expand_unique_f32", "b200emb_expand_unique_f64",
    "b200emb_gather_sorted_f32", "b200emb_gather_sorted_f64",
    "b200emb_lookup_table_scatter_f32", "b200emb_merge_add_scatter_f32", "b200emb_segment_scatter_i64",
    "b200emb_split_ids", "b200emb_merge_ids_f32", "b200emb_merge_ids_f64",
    "b200emb_split_selected_rows_f32", "b200emb_split_selected_rows_f64", "b200emb_extract_rows",
    "b200emb_sparse_decay_f32", "b200emb_sparse_decay_f64",
    "b200emb_peer_alloc", "b200emb_peer_open", "b200emb_peer_close", "b200emb_peer_free",
    "b200emb_peer_barrier", "b200emb_exchange_maps", "b200emb_peer_allgather_i64",
    "b200emb_shard_create", "b200emb_shard_export", "b200emb_shard_connect", "b200emb_shard_block",
    "b200emb_shard_connect_ptrs", "b200emb_shard_prepare", "b200emb_shard_lookup", "b200emb_shard_update",
    "b200emb_shard_lookup_phases", "b200emb_shard_update_phases", "b200emb_shard_status_ptr",
    "b200emb_shard_info", "b200emb_shard_debug_ptrs", "b200emb_shard_destroy",
]
STATUS_BARRIER_TIMEOUT = 4
STATUS_SHARD_OVERFLOW = 8

_lib = None


class B200EmbError(RuntimeError):
    pass


def lib():
    """Load libb200emb.so; raise loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise B200EmbError(
                "libb200emb.so not built (run __graft_entry__.build()); there is no CPU fallback")
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.b200emb_workspace_bytes.restype = ctypes.c_size_t
        _lib.b200emb_workspace_bytes.argtypes = [ctypes.c_int64] * 3
        _lib.b200emb_shard_block.restype = ctypes.c_void_p
        _lib.b200emb_shard_status_ptr.restype = ctypes.c_void_p
    return _lib


def _check(rc, what):
    if rc != 0:
        extra = ""
        if rc == -3:
            extra = " (cudaError %d)" % lib().b200emb_last_cuda_error()
        raise B200EmbError("%s failed: %s%s" % (what, _ERRORS.get(rc, rc), extra))


def _ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _stream(stream=None):
    # the raw handle of torch's current stream; torch.cuda.current_stream() builds a Stream object and
    # costs ~15 us per call, which adds up over the ~20 launches of a sharded step
    if stream is not None:
        return ctypes.c_void_p(stream if isinstance(stream, int) else stream.cuda_stream)
    return ctypes.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


def _sfx(dtype):
    if dtype == torch.float32:
        return "f32"
    if dtype == torch.float64:
        return "f64"
    raise TypeError("tables are float32 or float64 (lookup_table_op.cu:175-180), got %s" % dtype)


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B200EmbError("b200emb ops need CUDA tensors; there is no CPU fallback")
        if t is not None and not t.is_contiguous():
            raise B200EmbError("b200emb ops need contiguous tensors")


def radix_passes(n, height):
    """Radix passes of the sort plan (sort.cu digit_plan): digits of up to 10 bits."""
    bits = 1
    while (1 << bits) <= int(height):
        bits += 1
    return (bits + 9) // 10


HEAD_SCAN_KERNELS = 2   # k_head_count + k_head_write


def plan_kernel_count(n, height):
    """Kernels one sort plan launches: prepare + radix passes, plus one column scan per pass above
    64 tiles of 2048 ids (1 with B200EMB_CLUSTER_SORT set).  The run-length head scan that follows
    a plan (inside plan_unique_rows / the first consumer) is HEAD_SCAN_KERNELS more."""
    if 0 < n <= 16 * 8704 and "B200EMB_CLUSTER_SORT" in os.environ:
        return 1
    passes = radix_passes(n, height)
    return 1 + passes * (1 if (n + 2047) // 2048 <= 64 else 2)


def workspace_bytes(n, d, elem_size=4):
    return int(lib().b200emb_workspace_bytes(int(n), int(d), int(elem_size)))


def alloc_workspace(n, d, elem_size=4, device="cuda"):
    nbytes = max(workspace_bytes(n, d, elem_size), 256)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def alloc_workspace_upto(n_max, d, elem_size=4, device="cuda"):
    """Workspace valid for ANY n <= n_max.  The sort uses 1024-key tiles up to 1 212 416 ids and 2048-key
    tiles above (sort.cu sort_tile_keys), so the requirement is not monotonic across that boundary."""
    nbytes = max(workspace_bytes(n_max, d, elem_size), workspace_bytes(min(n_max, 148 * 4 * 2 * 1024), d, elem_size), 256)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def new_status(device="cuda"):
    return torch.zeros(1, dtype=torch.int32, device=device)


def lookup_table(table, ids, padding_idx=-1, out=None, status=None):
    """Out[k,:] = W[ids[k],:] (padding rows zero). ids: int64, any shape; returns [ids.numel(), D]."""
    _need_cuda(table, ids, out, status)
    assert ids.dtype == torch.int64
    v, d = table.shape
    n = ids.numel()
    if out is None:
        out = torch.empty((n, d), dtype=table.dtype, device=table.device)
    fn = getattr(lib(), "b200emb_lookup_table_" + _sfx(table.dtype))
    _check(fn(_ptr(table), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(ids), ctypes.c_int64(n),
              ctypes.c_int64(padding_idx), _ptr(out), _ptr(status), _stream()), "lookup_table")
    return out


def lookup_table_grad_sparse(ids, d_out, rows=None, value=None):
    """SelectedRows grad: rows = ids verbatim, value = copy of dOut (pass rows=ids/value=d_out to alias)."""
    _need_cuda(ids, d_out, rows, value)
    n = ids.numel()
    d = d_out.shape[-1]
    if rows is None:
        rows = torch.empty(n, dtype=torch.int64, device=ids.device)
    if value is None:
        value = torch.empty((n, d), dtype=d_out.dtype, device=d_out.device)
    fn = getattr(lib(), "b200emb_lookup_table_grad_sparse_" + _sfx(d_out.dtype))
    _check(fn(_ptr(ids), _ptr(d_out), ctypes.c_int64(n), ctypes.c_int64(d), _ptr(rows), _ptr(value),
              _stream()), "lookup_table_grad")
    return rows, value


def sum_selected_rows(inputs):
    """`sum` over SelectedRows: concat (rows, value) pairs, skipping empty ones (sum_op.h:71-145)."""
    total = sum(int(r.numel()) for r, _ in inputs)
    first = next((v for r, v in inputs if r.numel() > 0), inputs[-1][1])
    d = first.shape[1]
    out_rows = torch.empty(total, dtype=torch.int64, device=first.device)
    out_value = torch.empty((total, d), dtype=first.dtype, device=first.device)
    off = 0
    for r, v in inputs:
        n = int(r.numel())
        if n == 0:
            continue
        _need_cuda(r, v)
        _check(lib().b200emb_selected_rows_add_to(
            _ptr(r), _ptr(v), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(v.element_size()),
            _ptr(out_rows), _ptr(out_value), ctypes.c_int64(off), _stream()), "selected_rows_add_to")
        off += n
    return out_rows, out_value


def plan_rows(rows, height, workspace, status=None, unique=False, stream=None):
    """Sort plan of `rows` into the workspace.  unique=True also leaves the unique index there
    (plan_unique_rows with one shard and no outputs): consumers then take PLAN_READY | UNIQUE_READY."""
    _need_cuda(rows, workspace, status)
    if unique:
        _check(lib().b200emb_plan_unique_rows(
            _ptr(rows), ctypes.c_int64(rows.numel()), ctypes.c_int64(height), ctypes.c_int64(1), None, None,
            None, None, None, _ptr(workspace), ctypes.c_size_t(workspace.numel()), _ptr(status),
            _stream(stream)), "plan_unique_rows")
        return
    _check(lib().b200emb_plan_rows(_ptr(rows), ctypes.c_int64(rows.numel()), ctypes.c_int64(height),
                                   _ptr(workspace), ctypes.c_size_t(workspace.numel()), _ptr(status),
                                   _stream(stream)), "plan_rows")


def merge_add(rows, value, height, workspace=None, flags=0, status=None):
    """scatter::MergeAdd -> (out_rows[n], out_value[n,d], n_unique[1] int64); only the first U rows are meaningful."""
    _need_cuda(rows, value, workspace, status)
    n, d = value.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, value.element_size(), value.device)
    out_rows = torch.empty(n, dtype=torch.int64, device=value.device)
    out_value = torch.empty((n, d), dtype=value.dtype, device=value.device)
    n_unique = torch.zeros(1, dtype=torch.int64, device=value.device)
    fn = getattr(lib(), "b200emb_merge_add_" + _sfx(value.dtype))
    _check(fn(_ptr(rows), _ptr(value), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(height),
              _ptr(out_rows), _ptr(out_value), _ptr(n_unique), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), ctypes.c_uint(flags), _ptr(status), _stream()),
           "merge_add")
    return out_rows, out_value, n_unique


def selected_rows_add_to_tensor(rows, value, tensor, workspace=None, flags=0, status=None):
    _need_cuda(rows, value, tensor, workspace, status)
    n, d = value.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, value.element_size(), value.device)
    fn = getattr(lib(), "b200emb_selected_rows_add_to_tensor_" + _sfx(value.dtype))
    _check(fn(_ptr(rows), _ptr(value), ctypes.c_int64(n), ctypes.c_int64(d),
              ctypes.c_int64(tensor.shape[0]), _ptr(tensor), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), ctypes.c_uint(flags), _ptr(status), _stream()),
           "selected_rows_add_to_tensor")
    return tensor


def sparse_sgd(param, rows, grad, lr, workspace=None, flags=0, status=None):
    """In-place sparse SGD. lr: 1-element device tensor (the LearningRate var)."""
    _need_cuda(param, rows, grad, lr, workspace, status)
    n = rows.numel()
    v, d = param.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, param.element_size(), param.device)
    fn = getattr(lib(), "b200emb_sparse_sgd_" + _sfx(param.dtype))
    _check(fn(_ptr(param), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(rows), _ptr(grad),
              ctypes.c_int64(n), _ptr(lr), _ptr(workspace), ctypes.c_size_t(workspace.numel()),
              ctypes.c_uint(flags), _ptr(status), _stream()), "sparse_sgd")
    return param


def sparse_adam(param, moment1, moment2, rows, grad, lr, beta1_pow, beta2_pow, beta1=0.9, beta2=0.999,
                epsilon=1e-8, workspace=None, flags=0, status=None):
    """In-place sparse (lazy) Adam = MergeAdd fused with SparseAdamFunctor."""
    _need_cuda(param, moment1, moment2, rows, grad, lr, beta1_pow, beta2_pow, workspace, status)
    n = rows.numel()
    v, d = param.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, param.element_size(), param.device)
    fn = getattr(lib(), "b200emb_sparse_adam_" + _sfx(param.dtype))
    _check(fn(_ptr(param), _ptr(moment1), _ptr(moment2), ctypes.c_int64(v), ctypes.c_int64(d),
              _ptr(rows), _ptr(grad), ctypes.c_int64(n), _ptr(lr), _ptr(beta1_pow), _ptr(beta2_pow),
              ctypes.c_float(beta1), ctypes.c_float(beta2), ctypes.c_float(epsilon), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), ctypes.c_uint(flags), _ptr(status), _stream()),
           "sparse_adam")
    return param


def sparse_sgd_noalloc(param, rows, grad, lr, flags=0, status=None):
    """In-place sparse SGD (float32) for callers that have no workspace to give and must not
    allocate: the library's own fixed 8 MiB device array is the scratch.  Raises B200EmbError
    when workspace_bytes(n, d) exceeds NOALLOC_SCRATCH_BYTES.  Calls on one device share that
    array, so keep them on one stream (or otherwise serialised)."""
    _need_cuda(param, rows, grad, lr, status)
    if param.dtype != torch.float32:
        raise TypeError("sparse_sgd_noalloc is float32 only")
    n = rows.numel()
    v, d = param.shape
    _check(lib().b200emb_sparse_sgd_noalloc_f32(
        _ptr(param), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(rows), _ptr(grad), ctypes.c_int64(n),
        _ptr(lr), ctypes.c_uint(flags), _ptr(status), _stream()), "sparse_sgd_noalloc")
    return param


def plan_height(height, num_shards):
    """Height the plan is built for: num_shards * ceil(height / num_shards) (= height for one shard)."""
    if num_shards <= 1:
        return int(height)
    return int(num_shards) * ((int(height) + num_shards - 1) // num_shards)


def plan_unique_rows(rows, height, num_shards, workspace, status=None, want_elem_bounds=False, out=None,
                     stream=None):
    """Sort rows by (owner = id % num_shards, local row = id // num_shards) and de-duplicate.

    Returns (unique_keys[n], n_unique[1], shard_bounds[num_shards+1], elem_bounds or None, local_rows[n]);
    only the first U entries of unique_keys / local_rows are meaningful.  The workspace keeps the plan.
    `out` = the same five tensors preallocated by the caller (elem_bounds may be None); `stream` = the
    stream to launch on (default: torch's current stream)."""
    _need_cuda(rows, workspace, status)
    n = rows.numel()
    if out is None:
        dev = rows.device
        ukeys = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        local = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        n_unique = torch.zeros(1, dtype=torch.int64, device=dev)
        bounds = torch.zeros(num_shards + 1, dtype=torch.int64, device=dev)
        ebounds = torch.zeros(num_shards + 1, dtype=torch.int64, device=dev) if want_elem_bounds else None
    else:
        ukeys, n_unique, bounds, ebounds, local = out
        if not want_elem_bounds:
            ebounds = None
    _check(lib().b200emb_plan_unique_rows(
        _ptr(rows), ctypes.c_int64(n), ctypes.c_int64(height), ctypes.c_int64(num_shards), _ptr(ukeys),
        _ptr(n_unique), _ptr(bounds), _ptr(ebounds), _ptr(local), _ptr(workspace),
        ctypes.c_size_t(workspace.numel()), _ptr(status), _stream(stream)), "plan_unique_rows")
    return ukeys, n_unique, bounds, ebounds, local


def _permute(name, src, n, d, plan_h, workspace, out):
    _need_cuda(src, workspace, out)
    fn = getattr(lib(), name + _sfx(src.dtype))
    _check(fn(_ptr(src), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(plan_h), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), _ptr(out), _stream()), name)
    return out


def expand_unique(unique_values, n, plan_h, workspace, out=None):
    """out[i,:] = unique_values[unique index of rows[i], :] for the plan held by `workspace`."""
    d = unique_values.shape[1]
    if out is None:
        out = torch.empty((n, d), dtype=unique_values.dtype, device=unique_values.device)
    return _permute("b200emb_expand_unique_", unique_values, n, d, plan_h, workspace, out)


def gather_sorted(value, plan_h, workspace, out=None):
    """Rows of value[n,d] reordered into the plan's (owner, local row, position) order."""
    n, d = value.shape
    if out is None:
        out = torch.empty_like(value)
    return _permute("b200emb_gather_sorted_", value, n, d, plan_h, workspace, out)


# ---------------------------------------------------------------- peer-store (fused exchange) helpers
class PeerBuffer:
    """A device buffer that other ranks' GPUs can store into (CUDA IPC).  Set-up only."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _check(lib().b200emb_peer_alloc(ctypes.c_size_t(self.nbytes), ctypes.byref(ptr), handle), "peer_alloc")
        self.ptr = int(ptr.value)
        self.handle = bytes(handle)
        self.opened = {}

    def tensor(self, shape, dtype):
        """torch view of the first prod(shape) elements (no ownership)."""
        n = 1
        for x in shape:
            n *= int(x)
        base = self.opened.get(dtype)
        if base is None:          # one full-size view per dtype, sliced per call
            item = torch.empty(0, dtype=dtype).element_size()
            typestr = {torch.float32: "<f4", torch.float64: "<f8", torch.int64: "<i8", torch.uint8: "|u1"}[dtype]

            class _Iface:
                pass
            o = _Iface()
            o.__cuda_array_interface__ = {"shape": (self.nbytes // item,), "typestr": typestr,
                                          "data": (self.ptr, False), "version": 2}
            base = self.opened[dtype] = torch.as_tensor(o, device="cuda")
        assert n <= base.numel()
        return base[:n].view(tuple(int(x) for x in shape))

    def free(self):
        if self.ptr:
            lib().b200emb_peer_free(ctypes.c_void_p(self.ptr))
            self.ptr = 0


def peer_barrier(flag_ptrs, world, rank, epoch, status=None):
    """Stream-ordered barrier over peer memory; flag_ptrs = device int64[world] of the ranks' flag arrays."""
    _need_cuda(flag_ptrs, status)
    _check(lib().b200emb_peer_barrier(_ptr(flag_ptrs), ctypes.c_int(world), ctypes.c_int(rank),
                                      ctypes.c_uint64(epoch), _ptr(status), _stream()), "peer_barrier")


def peer_allgather_i64(row, table_ptrs, local_table, flag_ptrs, world, rank, epoch, host_copy=None,
                       status=None, stream=None):
    """All-gather of one int64 row per rank over peer memory (see b200emb_peer_allgather_i64); `host_copy` = a
    pinned host int64 tensor of world*row.numel() elements that receives the complete table."""
    _check(lib().b200emb_peer_allgather_i64(
        _ptr(row), ctypes.c_int(row.numel()), _ptr(table_ptrs), _ptr(local_table), _ptr(flag_ptrs),
        ctypes.c_int(world), ctypes.c_int(rank), ctypes.c_uint64(epoch), _ptr(host_copy), _ptr(status),
        _stream(stream)), "peer_allgather_i64")


def exchange_maps(split_table, rank, peer_i_ptrs, peer_u_ptrs, peer_g_ptrs, row_bytes, out=None, stream=None):
    """Scatter maps of the fused exchange from the all-gathered split table (device int64 [world, >=world+1]):
    out[0,:world] id destinations, out[1,:world+1] owner-gather segment starts, out[2,:world] owner-gather
    destinations, out[3,:world] gradient destinations (see b200emb_exchange_maps)."""
    _need_cuda(split_table, peer_i_ptrs, peer_u_ptrs, peer_g_ptrs, out)
    world, cols = split_table.shape
    if out is None:
        out = torch.empty((4, world + 1), dtype=torch.int64, device=split_table.device)
    _check(lib().b200emb_exchange_maps(_ptr(split_table), ctypes.c_int(world), ctypes.c_int(cols), ctypes.c_int(rank),
                                       _ptr(peer_i_ptrs), _ptr(peer_u_ptrs), _ptr(peer_g_ptrs),
                                       ctypes.c_int64(row_bytes), _ptr(out), _stream(stream)), "exchange_maps")
    return out


def peer_open(handle_bytes):
    ptr = ctypes.c_void_p()
    buf = (ctypes.c_ubyte * 64).from_buffer_copy(handle_bytes)
    _check(lib().b200emb_peer_open(buf, ctypes.byref(ptr)), "peer_open")
    return int(ptr.value)


def lookup_table_scatter(table, ids, seg_start, seg_dst, status=None, padding_idx=-1):
    """Gather rows table[ids[k]] and store row k of segment g at seg_dst[g] + (k - seg_start[g]) * d.
    seg_start: int64 device tensor [G+1]; seg_dst: int64 device tensor [G] of (possibly peer) addresses."""
    _need_cuda(table, ids, seg_start, seg_dst, status)
    v, d = table.shape
    _check(lib().b200emb_lookup_table_scatter_f32(
        _ptr(table), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(ids), ctypes.c_int64(ids.numel()),
        ctypes.c_int64(padding_idx), _ptr(seg_start), _ptr(seg_dst), ctypes.c_int64(seg_dst.numel()),
        _ptr(status), _stream()), "lookup_table_scatter")


def merge_add_scatter(rows, value, height, shard_bounds, shard_dst, workspace, flags=0, status=None):
    """MergeAdd whose merged row u is stored at shard_dst[g] + (u - shard_bounds[g]) * d (peer stores)."""
    _need_cuda(rows, value, shard_bounds, shard_dst, workspace, status)
    n, d = value.shape
    _check(lib().b200emb_merge_add_scatter_f32(
        _ptr(rows), _ptr(value), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(height),
        _ptr(shard_bounds), _ptr(shard_dst), ctypes.c_int64(shard_dst.numel()), None, _ptr(workspace),
        ctypes.c_size_t(workspace.numel()), ctypes.c_uint(flags), _ptr(status), _stream()), "merge_add_scatter")


def segment_scatter_i64(src, n_upper, bounds, dst):
    """dst[g][u - bounds[g]] = src[u] for every u < bounds[-1] (destinations may be peer GPUs)."""
    _need_cuda(src, bounds, dst)
    _check(lib().b200emb_segment_scatter_i64(_ptr(src), ctypes.c_int64(n_upper), _ptr(bounds), _ptr(dst),
                                             ctypes.c_int64(dst.numel()), _stream()), "segment_scatter_i64")


def sparse_adagrad(param, moment, rows, grad, lr, epsilon=1e-6, workspace=None, flags=0, status=None):
    """In-place sparse Adagrad = MergeAdd fused with the moment/parameter row update (adagrad_op.cc:108-141)."""
    _need_cuda(param, moment, rows, grad, lr, workspace, status)
    n = rows.numel()
    v, d = param.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, param.element_size(), param.device)
    fn = getattr(lib(), "b200emb_sparse_adagrad_" + _sfx(param.dtype))
    _check(fn(_ptr(param), _ptr(moment), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(rows), _ptr(grad),
              ctypes.c_int64(n), _ptr(lr), ctypes.c_float(epsilon), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), ctypes.c_uint(flags), _ptr(status), _stream()),
           "sparse_adagrad")
    return param


def split_ids(ids, height, shard_num, workspace, status=None):
    """Reference split_ids on a LoDTensor of ids: (out_ids[n] grouped by id % shard_num in input order,
    bounds[shard_num+1]).  The workspace keeps the permutation for merge_ids."""
    _need_cuda(ids, workspace, status)
    n = ids.numel()
    out = torch.empty(max(n, 1), dtype=torch.int64, device=ids.device)
    bounds = torch.zeros(shard_num + 1, dtype=torch.int64, device=ids.device)
    _check(lib().b200emb_split_ids(_ptr(ids), ctypes.c_int64(n), ctypes.c_int64(height), ctypes.c_int64(shard_num),
                                   _ptr(out), _ptr(bounds), _ptr(workspace), ctypes.c_size_t(workspace.numel()),
                                   _ptr(status), _stream()), "split_ids")
    return out[:n], bounds


def split_selected_rows(rows, value, height_sections, workspace=None):
    """Reference split_selected_rows: range split by height_sections (host list).  Returns
    (out_rows[n] rebased to the section start, out_value[n,d], bounds[len(height_sections)+1]); section s
    owns output positions [bounds[s], bounds[s+1])."""
    _need_cuda(rows, value, workspace)
    n, d = value.shape
    if workspace is None:
        workspace = alloc_workspace(n, d, value.element_size(), value.device)
    hs = (ctypes.c_int64 * len(height_sections))(*[int(h) for h in height_sections])
    out_rows = torch.empty(max(n, 1), dtype=torch.int64, device=value.device)
    out_value = torch.empty((n, d), dtype=value.dtype, device=value.device)
    bounds = torch.zeros(len(height_sections) + 1, dtype=torch.int64, device=value.device)
    fn = getattr(lib(), "b200emb_split_selected_rows_" + _sfx(value.dtype))
    _check(fn(_ptr(rows), _ptr(value), ctypes.c_int64(n), ctypes.c_int64(d), hs,
              ctypes.c_int64(len(height_sections)), _ptr(out_rows), _ptr(out_value), _ptr(bounds),
              _ptr(workspace), ctypes.c_size_t(workspace.numel()), _stream()), "split_selected_rows")
    return out_rows[:n], out_value, bounds


def extract_rows(rows, out=None):
    """Reference extract_rows: the SelectedRows' row ids as an int64 tensor [n, 1]."""
    _need_cuda(rows, out)
    n = rows.numel()
    if out is None:
        out = torch.empty((n, 1), dtype=torch.int64, device=rows.device)
    _check(lib().b200emb_extract_rows(_ptr(rows), ctypes.c_int64(n), _ptr(out), _stream()), "extract_rows")
    return out


DECAY_L2, DECAY_L1 = 0, 1


def sparse_decay(table, rows, coeff, mode=DECAY_L2, addend=None, out=None, status=None):
    """The regularizers' SelectedRows chain in one kernel: out[i,:] = addend[i,:] + coeff*f(table[rows[i],:])
    (f = identity for L2, sign for L1); addend=None gives the decay term alone."""
    _need_cuda(table, rows, addend, out, status)
    v, d = table.shape
    n = rows.numel()
    if out is None:
        out = torch.empty((n, d), dtype=table.dtype, device=table.device)
    ct = ctypes.c_float if table.dtype == torch.float32 else ctypes.c_double
    fn = getattr(lib(), "b200emb_sparse_decay_" + _sfx(table.dtype))
    _check(fn(_ptr(table), ctypes.c_int64(v), ctypes.c_int64(d), _ptr(rows), ctypes.c_int64(n), ct(coeff),
              ctypes.c_int(mode), _ptr(addend), _ptr(out), _ptr(status), _stream()), "sparse_decay")
    return out


def merge_ids(x, shard_num, workspace, out=None):
    """Reference merge_ids: rows of the shards (concatenated shard after shard) back to the order of the ids
    that split_ids was run on with this workspace."""
    _need_cuda(x, workspace, out)
    n, d = x.shape
    if out is None:
        out = torch.empty_like(x)
    fn = getattr(lib(), "b200emb_merge_ids_" + _sfx(x.dtype))
    _check(fn(_ptr(x), ctypes.c_int64(n), ctypes.c_int64(d), ctypes.c_int64(shard_num), _ptr(workspace),
              ctypes.c_size_t(workspace.numel()), _ptr(out), _stream()), "merge_ids")
    return out


# ---------------------------------------------------------------- row-sharded step engine (b200emb_shard_*)
def _view(ptr, n, dtype, device):
    """torch view of n elements of device memory owned by the library."""
    typestr = {torch.float32: "<f4", torch.int64: "<i8", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]

    class _Iface:
        pass
    o = _Iface()
    o.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(o, device=device)


class ShardEngine:
    """One rank of the row-sharded table (include/b200emb.h, b200emb_shard_*): owner(id) = id % world.
    `shard` [ceil(height/world), d] float32 (and moments for Adam) stay owned by the caller.  After
    construction exchange `handle` between the ranks and call connect(all_handles); ranks hosted in one
    process use connect_local(engines) instead."""

    SGD, ADAM = 0, 1

    def __init__(self, world, rank, height, d, shard, max_ids, max_recv, optimizer="sgd", m1=None, m2=None):
        _need_cuda(shard, m1, m2)
        assert shard.dtype == torch.float32 and shard.is_contiguous()
        self.world, self.rank, self.height, self.d = int(world), int(rank), int(height), int(d)
        self.device = shard.device
        self.shard, self.m1, self.m2 = shard, m1, m2
        self.h = ctypes.c_void_p()
        opt = self.ADAM if optimizer == "adam" else self.SGD
        with torch.cuda.device(self.device):
            _check(lib().b200emb_shard_create(
                ctypes.c_int(world), ctypes.c_int(rank), ctypes.c_int64(height), ctypes.c_int64(d), ctypes.c_int(opt),
                _ptr(shard), _ptr(m1), _ptr(m2), ctypes.c_int64(max_ids), ctypes.c_int64(max_recv),
                ctypes.byref(self.h)), "shard_create")
            hb = (ctypes.c_ubyte * 64)()
            _check(lib().b200emb_shard_export(self.h, hb), "shard_export")
        self.handle = bytes(hb)
        self.block = int(lib().b200emb_shard_block(self.h))
        info = (ctypes.c_int64 * 8)()
        _check(lib().b200emb_shard_info(self.h, info), "shard_info")
        self.hs, self.plan_h, self.cap_ids, self.cap_src, self.cap_recv, self.block_bytes = (int(info[i]) for i in range(2, 8))
        self.status = _view(lib().b200emb_shard_status_ptr(self.h), 1, torch.int32, self.device)
        self._n = [0] * 4
        self._keep = [None] * 4

    def connect(self, handles):
        buf = (ctypes.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(handles))
        with torch.cuda.device(self.device):
            _check(lib().b200emb_shard_connect(self.h, buf), "shard_connect")

    @staticmethod
    def connect_local(engines):
        """Ranks living in this process (one or several GPUs with peer access): hand every engine all blocks."""
        blocks = (ctypes.c_void_p * len(engines))(*[e.block for e in engines])
        for e in engines:
            with torch.cuda.device(e.device):
                _check(lib().b200emb_shard_connect_ptrs(e.h, blocks), "shard_connect_ptrs")

    def prepare(self, slot, ids, stream=None):
        _need_cuda(ids)
        assert ids.dtype == torch.int64 and ids.is_contiguous()
        self._n[slot] = ids.numel()
        self._keep[slot] = ids
        _check(lib().b200emb_shard_prepare(self.h, ctypes.c_int(slot), _ptr(ids), ctypes.c_int64(ids.numel()),
                                           _stream(stream)), "shard_prepare")

    def lookup(self, slot, out=None, stream=None, phases=None):
        if out is None:
            out = torch.empty((self._n[slot], self.d), dtype=torch.float32, device=self.device)
        if phases is None:
            _check(lib().b200emb_shard_lookup(self.h, ctypes.c_int(slot), _ptr(out), _stream(stream)), "shard_lookup")
        else:
            _check(lib().b200emb_shard_lookup_phases(self.h, ctypes.c_int(slot), _ptr(out), _stream(stream),
                                                     ctypes.c_int(phases[0]), ctypes.c_int(phases[1])), "shard_lookup_phases")
        return out

    def update(self, slot, d_out, lr, b1p=None, b2p=None, beta1=0.9, beta2=0.999, eps=1e-8, stream=None, phases=None):
        _need_cuda(d_out, lr, b1p, b2p)
        args = (self.h, ctypes.c_int(slot), _ptr(d_out), _ptr(lr), _ptr(b1p), _ptr(b2p), ctypes.c_float(beta1),
                ctypes.c_float(beta2), ctypes.c_float(eps), _stream(stream))
        if phases is None:
            _check(lib().b200emb_shard_update(*args), "shard_update")
        else:
            _check(lib().b200emb_shard_update_phases(*args, ctypes.c_int(phases[0]), ctypes.c_int(phases[1])),
                   "shard_update_phases")

    def debug_views(self, slot):
        """(split points [world+1], unique count [1], owner segment starts [world+1], step counter) as device views."""
        out = (ctypes.c_void_p * 6)()
        _check(lib().b200emb_shard_debug_ptrs(self.h, ctypes.c_int(slot), out), "shard_debug_ptrs")
        w = self.world
        return (_view(out[0], w + 1, torch.int64, self.device), _view(out[1], 1, torch.int64, self.device),
                _view(out[2], w + 1, torch.int64, self.device), _view(out[4], 1, torch.int64, self.device))

    def check_status(self):
        """Raise if a kernel of the sharded step flagged a time-out or a capacity overflow (reads the device word)."""
        st = int(self.status.item())
        if st & STATUS_BARRIER_TIMEOUT:
            raise B200EmbError("sharded step: a rank did not arrive within 20 s (status %d)" % st)
        if st & STATUS_SHARD_OVERFLOW:
            raise B200EmbError("sharded step: more rows than the capacities given at create() (status %d)" % st)
        return st

    def destroy(self):
        if self.h:
            lib().b200emb_shard_destroy(self.h)
            self.h = ctypes.c_void_p()
