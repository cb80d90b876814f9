"""Validates the restated oracle against the reference's OWN CPU kernels (oracle/_ref, compiled
unmodified from /root/reference with the repo's compat headers) — bit for bit on random and golden
inputs.  Runs wherever oracle/_ref/libref_cpu.so exists (it is built in the container that mounts
the reference and travels to the GPU box as a built artefact); skipped otherwise."""
import numpy as np
import pytest

import oracle
from oracle import ref_cpu

pytestmark = pytest.mark.skipif(not ref_cpu.available(), reason="oracle/_ref not built (no /root/reference here)")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32 if a.dtype == np.float32 else np.uint64)


def case(seed, n, v, d, dtype):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, v, size=n).astype(np.int64)
    rows[: n // 3] = rows[0]            # a hot row
    value = rng.standard_normal((n, d)).astype(dtype)
    return rng, rows, value


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,v,d", [(1, 3, 4), (50, 9, 7), (700, 64, 33), (3000, 1000, 64)])
def test_lookup_and_grad(dtype, n, v, d):
    rng, ids, dout = case(n, n, v, d, dtype)
    table = rng.standard_normal((v, d)).astype(dtype)
    for pad in (-1, int(ids[-1])):
        assert np.array_equal(bits(ref_cpu.lookup_table(table, ids, pad)), bits(oracle.lookup_table(table, ids, pad)))
    r1, v1 = ref_cpu.lookup_table_grad_sparse(ids, dout, v)
    r2, v2 = oracle.lookup_table_grad_sparse(ids, dout)
    assert np.array_equal(r1, r2) and np.array_equal(bits(v1), bits(v2))
    with pytest.raises(IndexError):
        ref_cpu.lookup_table(table, [v])
    with pytest.raises(IndexError):
        oracle.lookup_table(table, [v])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,v,d", [(1, 3, 4), (50, 9, 7), (700, 64, 33), (3000, 1000, 64)])
def test_merge_add_and_add_to_tensor(dtype, n, v, d):
    rng, rows, value = case(n + 1, n, v, d, dtype)
    r1, m1 = ref_cpu.merge_add(rows, value, v)
    for ref_form in (True, False):
        r2, m2 = oracle.merge_add(rows, value, reference_form=ref_form)
        assert np.array_equal(r1, r2) and np.array_equal(bits(m1), bits(m2))
    base = rng.standard_normal((v, d)).astype(dtype)
    t1, t2 = base.copy(), base.copy()
    ref_cpu.selected_rows_add_to_tensor(rows, value, t1)
    oracle.selected_rows_add_to_tensor(rows, value, t2)
    assert np.array_equal(bits(t1), bits(t2))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,v,d", [(1, 3, 4), (50, 9, 7), (700, 64, 33), (3000, 1000, 64)])
def test_sgd_and_adam(dtype, n, v, d):
    rng, rows, grad = case(n + 2, n, v, d, dtype)
    grad *= dtype(0.01)
    base = (rng.random((v, d)) * 0.1 - 0.05).astype(dtype)
    lr = np.array([0.01], dtype)
    p1, p2 = base.copy(), base.copy()
    ref_cpu.sgd_sparse(p1, rows, grad, lr)
    oracle.sgd_sparse(p2, rows, grad, lr)
    assert np.array_equal(bits(p1), bits(p2))
    m1a = (rng.standard_normal((v, d)) * 1e-3).astype(dtype)
    m2a = (rng.random((v, d)) * 1e-4).astype(dtype)
    b1p, b2p = np.array([0.9 ** 3], dtype), np.array([0.999 ** 3], dtype)
    a = [base.copy(), m1a.copy(), m2a.copy()]
    b = [base.copy(), m1a.copy(), m2a.copy()]
    ref_cpu.adam_sparse(a[0], a[1], a[2], rows, grad, lr, b1p, b2p)
    oracle.adam_sparse(b[0], b[1], b[2], rows, grad, lr, b1p, b2p, reference_form=True)
    for x, y in zip(a, b):
        assert np.array_equal(bits(x), bits(y))


def test_sum_selected_rows_with_empty_inputs(goldens):
    g = goldens["G7_sum_selected_rows"]
    full = np.ones((7, g["row_numel"]), np.float32) * np.arange(7, dtype=np.float32)[:, None]
    empty = np.ones((0, g["row_numel"]), np.float32)
    for c in g["cases"]:
        ins = [(g["rows"], full) if has else ([], empty) for has in c]
        r1, v1 = ref_cpu.sum_selected_rows(ins, height=7)
        r2, v2 = oracle.sum_selected_rows(ins)
        assert len(r1) == 7 * sum(c)
        assert np.array_equal(r1, r2) and np.array_equal(v1, v2)


def test_goldens_through_the_reference_kernels(goldens):
    # the reference's own test constants hold for its own kernels as driven here (sanity of the driver)
    g = goldens["G3_sparse_sgd"]
    grad = np.ones((3, g["row_numel"]), np.float32)
    for r, c, v in g["grad_set"]:
        grad[r, c] = v
    p = np.full((g["height"], g["row_numel"]), g["param_fill"], np.float32)
    ref_cpu.sgd_sparse(p, g["rows"], grad, [g["lr"]])
    for r, c, v in g["expect"]:
        assert p[r, c] == pytest.approx(v, abs=1e-7)
    g = goldens["G5_sparse_adagrad"]
    grad = np.ones((4, g["row_numel"]), np.float32)
    mr, mv = ref_cpu.merge_add(g["rows"], grad, g["height"])
    assert mr.tolist() == [0, 4, 7] and (mv[1] == 2.0).all()
