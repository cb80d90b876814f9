"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/b200emb.h
declares, and its pure-host entry points behave.  No kernel is launched."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200emb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200emb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(pkg):
    pkg.build_native()
    lib = pkg.capi.lib()
    declared = _declared_symbols()
    assert len(declared) >= 17
    for name in declared:
        assert hasattr(lib, name), "libb200emb.so does not export %s" % name
    assert set(declared) == set(pkg.capi.EXPORTS)
    assert lib.b200emb_abi_version() == 1


def test_workspace_bytes_is_monotonic_and_aligned(pkg):
    pkg.build_native()
    w = pkg.capi.workspace_bytes
    assert w(0, 64) > 0 and w(0, 64) % 256 == 0
    assert w(106496, 64) < w(655360, 128) < w(6553600, 256)
    assert w(1000, 64, 8) > w(1000, 64, 4)
    assert w(-1, 64) == 0 and w(10, 0) == 0 and w(10, 4, 3) == 0


def test_workspace_for_a_range_of_batch_sizes(pkg, monkeypatch):
    """alloc_workspace_upto(n_max) must cover every n <= n_max although the sort switches from 1024- to
    2048-key tiles above 1 212 416 ids, which makes the requirement drop there (host-only arithmetic)."""
    pkg.build_native()
    capi = pkg.capi
    edge = 148 * 4 * 2 * 1024
    assert capi.workspace_bytes(edge, 64) > capi.workspace_bytes(edge + 1, 64)      # the non-monotonic spot
    sizes = {}
    import torch
    monkeypatch.setattr(torch, "empty", lambda nbytes, **kw: sizes.setdefault("n", nbytes))
    for n_max in (1000, 131072, 131073, edge - 5, edge, edge + 1, 1703936, 3 * edge):
        sizes.clear()
        capi.alloc_workspace_upto(n_max, 64, 4, "cpu")
        probe = {1, n_max // 3, n_max // 2, n_max - 1, n_max, min(n_max, 131072), min(n_max, 131073), min(n_max, edge),
                 min(n_max, edge + 1)}
        assert all(capi.workspace_bytes(max(n, 1), 64) <= sizes["n"] for n in probe), n_max


def test_argument_errors_without_a_gpu(pkg):
    pkg.build_native()
    lib = pkg.capi.lib()
    null = ctypes.c_void_p(0)
    # negative sizes / zero width -> B200EMB_E_INVALID before anything touches the device
    assert lib.b200emb_lookup_table_f32(null, ctypes.c_int64(10), ctypes.c_int64(0), null,
                                        ctypes.c_int64(4), ctypes.c_int64(-1), null, null, null) == -1
    assert lib.b200emb_sparse_sgd_f32(null, ctypes.c_int64(10), ctypes.c_int64(4), null, null,
                                      ctypes.c_int64(-3), null, null, ctypes.c_size_t(0),
                                      ctypes.c_uint(0), null, null) == -1
    # beyond the documented limits -> B200EMB_E_UNSUPPORTED
    assert lib.b200emb_sparse_sgd_f32(null, ctypes.c_int64(1 << 33), ctypes.c_int64(4), null, null,
                                      ctypes.c_int64(8), null, null, ctypes.c_size_t(0),
                                      ctypes.c_uint(0), null, null) == -4
    # empty gradient is a legal no-op (sgd_op.h:56-58)
    assert lib.b200emb_sparse_sgd_f32(null, ctypes.c_int64(10), ctypes.c_int64(4), null, null,
                                      ctypes.c_int64(0), null, null, ctypes.c_size_t(0),
                                      ctypes.c_uint(0), null, null) == 0


def test_ops_refuse_cpu_tensors(pkg):
    import torch
    pkg.build_native()
    t = torch.zeros(4, 4)
    ids = torch.zeros(2, dtype=torch.int64)
    with pytest.raises(pkg.capi.B200EmbError):
        pkg.capi.lookup_table(t, ids)
