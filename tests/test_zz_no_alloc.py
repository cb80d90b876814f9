"""GPU test of b200emb_sparse_sgd_noalloc_f32 — the workspace-free sparse SGD for callers that
only hold an op's inputs and outputs and are not allowed to allocate device memory (thin FFI
shims, processes whose framework pool owns the GPU's memory).  It is called through the C ABI and
must reproduce b200emb_sparse_sgd_f32 on two calls in a row: the second call finds the first
call's leftovers in the library's built-in scratch array and must not be disturbed by them.

Collected last on purpose (file name), so a run that stops at the first failure has already
executed every other test; it also passes when run alone in a fresh process.
"""
import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FAST_RTOL = 1e-6   # the bound tests/test_gpu_parity.py uses for the fast mode


@pytest.fixture(scope="module")
def capi(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg.capi.lib()
    return pkg.capi


def _case(seed, n, v, d):
    rng = np.random.default_rng(seed)
    w = np.arange(1, v + 1, dtype=np.float64) ** -1.1
    cdf = np.cumsum(w) / w.sum()
    rows = ((np.searchsorted(cdf, rng.random(n)) * 2654435761 + 12345) % v).astype(np.int64)
    grad = (rng.standard_normal((n, d)) * 0.01).astype(np.float32)
    base = (rng.random((v, d)) * 0.1 - 0.05).astype(np.float32)
    return rows, grad, base


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("flags", [0, 1], ids=["fast", "deterministic"])
def test_noalloc_matches_workspace_entry_point_twice_in_a_row(capi, flags):
    lr = _dev(np.array([0.05], np.float32))
    # two different shapes back to back: the second call reuses the scratch the first one dirtied
    for seed, (n, v, d) in enumerate([(40000, 60000, 64), (9000, 5000, 128)]):
        rows, grad, base = _case(seed, n, v, d)
        assert capi.workspace_bytes(n, d, 4) <= capi.NOALLOC_SCRATCH_BYTES
        want = _dev(base)
        capi.sparse_sgd(want, _dev(rows), _dev(grad), lr, flags=flags)
        got = _dev(base)
        capi.sparse_sgd_noalloc(got, _dev(rows), _dev(grad), lr, flags=flags)
        torch.cuda.synchronize()
        g, w = got.cpu().numpy(), want.cpu().numpy()
        # same rows, same plan, same reduction tree: the two entry points agree bit for bit in BOTH modes
        assert np.array_equal(g.view(np.uint32), w.view(np.uint32))
        # ... and against the oracle: bit-exact in the reference-order mode, within the tolerance the other
        # GPU tests use for the fast mode (reordering bound against the fp64 evaluation)
        if flags == 1:
            ref = base.copy()
            oracle.sgd_sparse(ref, rows, grad, np.array([0.05], np.float32))
            assert np.array_equal(g.view(np.uint32), ref.view(np.uint32))
        urows, merged = oracle.merge_add(rows, grad, truth=True)
        _, l1 = oracle.merge_add(rows, np.abs(grad), truth=True)
        truth = base.astype(np.float64)
        truth[urows] -= 0.05 * merged
        bound = FAST_RTOL * np.abs(base).astype(np.float64)
        bound[urows] += FAST_RTOL * 0.05 * l1
        if flags == 0:
            assert (np.abs(g - truth) <= bound + 1e-30).all()
        assert not np.array_equal(g, base)   # the update really happened


def test_noalloc_rejects_requests_beyond_the_scratch(capi):
    n, v, d = 300000, 1000, 64
    assert capi.workspace_bytes(n, d, 4) > capi.NOALLOC_SCRATCH_BYTES
    param = torch.zeros((v, d), device="cuda")
    rows = torch.zeros(n, dtype=torch.int64, device="cuda")
    grad = torch.zeros((n, d), device="cuda")
    lr = torch.ones(1, device="cuda")
    with pytest.raises(capi.B200EmbError, match="built-in 8 MiB scratch"):
        capi.sparse_sgd_noalloc(param, rows, grad, lr)
    # an empty gradient is still a no-op, and a stale-plan flag is refused
    capi.sparse_sgd_noalloc(param, rows[:0], grad[:0], lr)
    with pytest.raises(capi.B200EmbError):
        capi.sparse_sgd_noalloc(param, rows[:10], grad[:10], lr, flags=capi.PLAN_READY)
