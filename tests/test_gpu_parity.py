"""GPU parity tests proper: the CUDA path, called through the C ABI (paddle-1_b200/capi.py is a
thin ctypes binding), against the CPU oracle on identical seeded inputs.

Bars (BASELINE.json north_star):
  * gathered rows, SelectedRows rows, merged row-index sets: bit-exact;
  * deterministic-order mode (B200EMB_DETERMINISTIC): merged values and updated parameters
    bit-exact against the reference CPU order;
  * fast mode: |got - want| <= 1e-6 * (sum of |addends| of that element) — i.e. 1e-6 relative to
    the magnitude of what was summed (a reordering cannot do better than that on cancelling
    sums) — and updated parameters within 1e-6 relative of |param| plus that bound.
"""
import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

DET = 1
FAST_RTOL = 1e-6


@pytest.fixture(scope="module")
def capi(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg.capi.lib()
    return pkg.capi


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t):
    return t.detach().cpu().numpy()


def bits(a):
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


def assert_bit_equal(got, want, what=""):
    assert got.shape == want.shape, what
    same = bits(np.ascontiguousarray(got)) == bits(np.ascontiguousarray(want))
    assert same.all(), "%s: %d of %d elements differ in bits" % (what, (~same).sum(), same.size)


def l1_of_segments(rows, value):
    """sum_i |value_i| per unique row, fp64 (the scale a reordered sum is accurate to)."""
    _, l1 = oracle.merge_add(rows, np.abs(value), truth=True)
    return l1


def make_case(rng, n, v, d, dtype=np.float32, dist="uniform"):
    if dist == "uniform":
        rows = rng.integers(0, v, size=n).astype(np.int64)
    elif dist == "zipf":
        w = np.arange(1, v + 1, dtype=np.float64) ** -1.1
        cdf = np.cumsum(w) / w.sum()
        rank = np.searchsorted(cdf, rng.random(n))
        rows = ((rank * 2654435761 + 12345) % v).astype(np.int64)
    elif dist == "same":
        rows = np.full(n, v // 2, dtype=np.int64)
    elif dist == "unique":
        rows = rng.permutation(v)[:n].astype(np.int64)
    value = rng.standard_normal((n, d)).astype(dtype)
    return rows, value


SHAPES = [  # (n, v, d, dist)
    (1, 5, 4, "uniform"), (3, 10, 12, "uniform"), (31, 7, 2, "uniform"), (257, 40, 31, "uniform"),
    (1000, 300, 10, "uniform"), (5000, 100000, 64, "zipf"), (5000, 4096, 32, "zipf"),
    (20000, 100000, 128, "zipf"), (6000, 900, 256, "zipf"), (3000, 17, 300, "uniform"),
    (4097, 1 << 20, 64, "same"), (9000, 9000, 64, "unique"), (2500, 70000, 520, "zipf"),
    (70000, 1 << 17, 8, "zipf"), (1500, 3000, 1028, "uniform"),
    # beyond the single-cluster sort's capacity (139 264 ids): multi-launch radix sort path
    (250000, 1 << 20, 16, "zipf"), (700000, 5000000, 8, "zipf"),
]


# ------------------------------------------------------------------ lookup_table (K1)
@pytest.mark.parametrize("d", [1, 2, 12, 31, 32, 64, 100, 128, 256, 300, 512, 1028])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_lookup_table_bit_exact(capi, d, dtype):
    rng = np.random.default_rng(d)
    v, n = 1000, 3333
    table = rng.standard_normal((v, d)).astype(dtype)
    ids = rng.integers(0, v, size=(n, 1)).astype(np.int64)
    for pad in (-1, int(ids[7, 0])):
        want = oracle.lookup_table(table, ids, pad)
        got = host(capi.lookup_table(dev(table), dev(ids), pad))
        assert_bit_equal(got, want, "lookup d=%d pad=%d" % (d, pad))


def test_lookup_table_golden_g6(capi, goldens):
    g = goldens["G6_lookup_table"]
    rng = np.random.default_rng(0)
    table = rng.random(g["table_shape"]).astype(np.float32)
    for shape in g["ids_shapes"]:
        ids = rng.integers(0, g["table_shape"][0], size=shape).astype(np.int64)
        got = host(capi.lookup_table(dev(table), dev(ids)))
        assert_bit_equal(got, table[ids.reshape(-1)], "G6")


def test_lookup_table_flags_out_of_range_ids(capi):
    table = np.ones((5, 8), np.float32)
    status = capi.new_status()
    out = host(capi.lookup_table(dev(table), dev(np.array([0, 5, -1, 4], np.int64)), status=status))
    assert int(status.item()) & 1
    assert (out[0] == 1).all() and (out[3] == 1).all() and (out[1] == 0).all() and (out[2] == 0).all()
    status.zero_()
    capi.lookup_table(dev(table), dev(np.array([0, 4], np.int64)), status=status)
    assert int(status.item()) == 0


def test_lookup_table_empty(capi):
    out = capi.lookup_table(dev(np.ones((5, 8), np.float32)), dev(np.zeros((0, 1), np.int64)))
    assert tuple(out.shape) == (0, 8)


def test_lookup_table_grad_sparse_and_sum(capi, goldens):
    rng = np.random.default_rng(3)
    ids = rng.integers(0, 50, size=(40, 1)).astype(np.int64)
    dout = rng.standard_normal((40, 12)).astype(np.float32)
    rows, value = capi.lookup_table_grad_sparse(dev(ids), dev(dout))
    wr, wv = oracle.lookup_table_grad_sparse(ids, dout)
    assert np.array_equal(host(rows), wr)
    assert_bit_equal(host(value), wv)
    # G7: sum over SelectedRows with empty inputs = concat of the non-empty ones
    g = goldens["G7_sum_selected_rows"]
    full = np.ones((7, g["row_numel"]), np.float32) * np.arange(7, dtype=np.float32)[:, None]
    for case in g["cases"]:
        ins = [(np.array(g["rows"] if has else [], np.int64),
                full if has else np.zeros((0, g["row_numel"]), np.float32)) for has in case]
        r, v = capi.sum_selected_rows([(dev(a), dev(b)) for a, b in ins])
        wr, wv = oracle.sum_selected_rows(ins)
        assert len(r) == 7 * sum(case)
        assert np.array_equal(host(r), wr) and np.array_equal(host(v), wv)


# ------------------------------------------------------------------ MergeAdd (K3)
@pytest.mark.parametrize("case", SHAPES, ids=lambda c: "n%d_v%d_d%d_%s" % c)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_merge_add(capi, case, dtype):
    n, v, d, dist = case
    rng = np.random.default_rng(n * 31 + d)
    rows, value = make_case(rng, n, v, d, dtype, dist)
    want_rows, want_val = oracle.merge_add(rows, value)
    # deterministic-order mode: bit-exact
    r, val, u = capi.merge_add(dev(rows), dev(value), v, flags=DET)
    u = int(u.item())
    assert u == len(want_rows)
    assert np.array_equal(host(r)[:u], want_rows)
    assert_bit_equal(host(val)[:u], want_val, "merge_add det")
    # fast mode: row set bit-exact, values within the reordering bound
    r, val, u2 = capi.merge_add(dev(rows), dev(value), v)
    assert int(u2.item()) == u and np.array_equal(host(r)[:u], want_rows)
    _, truth = oracle.merge_add(rows, value, truth=True)
    l1 = l1_of_segments(rows, value)
    tol = (FAST_RTOL if dtype == np.float32 else 1e-14) * l1 + 1e-30
    assert (np.abs(host(val)[:u].astype(np.float64) - truth) <= tol).all()
    # fast mode is run-to-run deterministic
    r3, val3, _ = capi.merge_add(dev(rows), dev(value), v)
    assert_bit_equal(host(val3)[:u], host(val)[:u], "merge_add fast repeat")


def test_merge_add_golden_g5_duplicate_rows(capi, goldens):
    g = goldens["G5_sparse_adagrad"]
    grad = np.ones((4, g["row_numel"]), np.float32)
    for r_, c_, v_ in g["grad_set"]:
        grad[r_, c_] = v_
    for flags in (0, DET):
        r, val, u = capi.merge_add(dev(np.array(g["rows"], np.int64)), dev(grad), g["height"], flags=flags)
        assert int(u.item()) == 3 and host(r)[:3].tolist() == [0, 4, 7]
        assert (host(val)[1] == 2.0).all() and host(val)[0, 0] == 2.0 and host(val)[2, 8] == 4.0


def test_merge_add_negative_zero(capi):
    rows = np.array([3, 3, 1], np.int64)
    value = np.full((3, 4), -0.0, np.float32)
    for flags in (0, DET):
        _, val, u = capi.merge_add(dev(rows), dev(value), 8, flags=flags)
        got = host(val)[:int(u.item())]
        want = oracle.merge_add(rows, value)[1]
        assert_bit_equal(got, want)


def test_merge_add_flags_bad_rows_and_skips_them(capi):
    rows = np.array([2, 9, 2, -4, 1], np.int64)
    value = np.ones((5, 4), np.float32)
    status = capi.new_status()
    r, val, u = capi.merge_add(dev(rows), dev(value), 5, status=status)
    assert int(status.item()) & 2
    assert int(u.item()) == 2 and host(r)[:2].tolist() == [1, 2]
    assert (host(val)[1] == 2.0).all()


def test_golden_g1_g2_concat_then_add_to_tensor(capi, goldens):
    g = goldens["G1_G2_selected_rows_add"]
    n = g["row_numel"]
    v1 = np.full((3, n), g["fill1"], np.float32)
    v2 = np.full((4, n), g["fill2"], np.float32)
    rows, value = capi.sum_selected_rows([(dev(np.array(g["rows1"], np.int64)), dev(v1)),
                                          (dev(np.array(g["rows2"], np.int64)), dev(v2))])
    assert host(rows).tolist() == g["expect_rows"]
    for r_, c_, v_ in g["expect_concat"]:
        assert host(value)[r_, c_] == v_
    for flags in (0, DET):
        t = dev(np.full((g["height"], n), g["tensor_fill"], np.float32))
        capi.selected_rows_add_to_tensor(rows, value, t, flags=flags)
        for r_, c_, v_ in g["expect_dense"]:
            assert host(t)[r_, c_] == v_


@pytest.mark.parametrize("case", SHAPES[3:10], ids=lambda c: "n%d_v%d_d%d_%s" % c)
def test_selected_rows_add_to_tensor(capi, case):
    n, v, d, dist = case
    rng = np.random.default_rng(n + d)
    rows, value = make_case(rng, n, v, d, np.float32, dist)
    base = rng.standard_normal((v, d)).astype(np.float32)
    want = base.copy()
    oracle.selected_rows_add_to_tensor(rows, value, want)
    t = dev(base)
    capi.selected_rows_add_to_tensor(dev(rows), dev(value), t, flags=DET)
    assert_bit_equal(host(t), want, "add_to_tensor det")
    t = dev(base)
    capi.selected_rows_add_to_tensor(dev(rows), dev(value), t)
    l1 = np.zeros((v, d))
    ur, l1u = oracle.merge_add(rows, np.abs(value), truth=True)
    l1[ur] = l1u
    assert (np.abs(host(t).astype(np.float64) - want) <= FAST_RTOL * (l1 + np.abs(base)) + 1e-30).all()


# ------------------------------------------------------------------ sparse SGD (K5a)
def test_sparse_sgd_golden_g3(capi, goldens):
    g = goldens["G3_sparse_sgd"]
    grad = np.ones((3, g["row_numel"]), np.float32)
    for r_, c_, v_ in g["grad_set"]:
        grad[r_, c_] = v_
    for flags in (0, DET):
        p = dev(np.full((g["height"], g["row_numel"]), g["param_fill"], np.float32))
        capi.sparse_sgd(p, dev(np.array(g["rows"], np.int64)), dev(grad), dev(np.array([g["lr"]], np.float32)),
                        flags=flags)
        for r_, c_, v_ in g["expect"]:
            assert host(p)[r_, c_] == pytest.approx(v_, abs=1e-6)


@pytest.mark.parametrize("case", SHAPES, ids=lambda c: "n%d_v%d_d%d_%s" % c)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_sparse_sgd(capi, case, dtype):
    n, v, d, dist = case
    rng = np.random.default_rng(n * 7 + d)
    rows, grad = make_case(rng, n, v, d, dtype, dist)
    grad *= dtype(0.01)
    base = (rng.random((v, d)) * 0.1 - 0.05).astype(dtype)
    lr = np.array([0.01], dtype)
    want = base.copy()
    oracle.sgd_sparse(want, rows, grad, lr)
    p = dev(base)
    capi.sparse_sgd(p, dev(rows), dev(grad), dev(lr), flags=DET)
    assert_bit_equal(host(p), want, "sgd det")
    p = dev(base)
    capi.sparse_sgd(p, dev(rows), dev(grad), dev(lr))
    l1 = np.zeros((v, d))
    ur, l1u = oracle.merge_add(rows, np.abs(grad), truth=True)
    l1[ur] = l1u
    rt = FAST_RTOL if dtype == np.float32 else 1e-14
    # held to the fp64 evaluation of the same update: on long duplicate runs the reference's own
    # sequential fp32 chain drifts further from it than 1e-6 relative
    truth = base.astype(np.float64)
    truth[ur] -= float(lr[0]) * oracle.merge_add(rows, grad, truth=True)[1]
    err = np.abs(host(p).astype(np.float64) - truth)
    assert (err <= rt * (np.abs(base) + float(lr[0]) * l1) + 1e-30).all()
    untouched = np.setdiff1d(np.arange(v), rows)
    assert_bit_equal(host(p)[untouched], base[untouched], "sgd untouched rows")


def test_sparse_sgd_empty_grad_is_noop(capi):
    base = np.ones((6, 8), np.float32)
    p = dev(base)
    capi.sparse_sgd(p, dev(np.zeros(0, np.int64)), dev(np.zeros((0, 8), np.float32)),
                    dev(np.array([0.5], np.float32)))
    assert np.array_equal(host(p), base)


def test_plan_ready_path_matches(capi):
    rng = np.random.default_rng(11)
    rows, grad = make_case(rng, 30000, 50000, 64, np.float32, "zipf")
    base = rng.standard_normal((50000, 64)).astype(np.float32)
    lr = dev(np.array([0.1], np.float32))
    ws = capi.alloc_workspace(30000, 64)
    for flags in (0, DET):
        p1 = dev(base)
        capi.sparse_sgd(p1, dev(rows), dev(grad), lr, workspace=ws, flags=flags)
        p2 = dev(base)
        drows = dev(rows)
        capi.plan_rows(drows, 50000, ws)
        capi.sparse_sgd(p2, drows, dev(grad), lr, workspace=ws, flags=flags | capi.PLAN_READY)
        assert_bit_equal(host(p2), host(p1), "plan-ready")


# ------------------------------------------------------------------ sparse Adam (K5b)
def test_sparse_adam_golden_g4(capi, goldens):
    g = goldens["G4_sparse_adam"]
    h, n = g["height"], g["row_numel"]
    grad = np.ones((3, n), np.float32)
    for r_, c_, v_ in g["grad_set"]:
        grad[r_, c_] = v_
    b1, b2, eps = g["beta1"], g["beta2"], g["epsilon"]
    b1p = np.array([b1 ** 10], np.float32)
    b2p = np.array([b2 ** 10], np.float32)
    lr = np.full((1,), g["lr"], np.float32)
    wp, wm1, wm2 = (np.full((h, n), 5.0, np.float32) for _ in range(3))
    oracle.adam_sparse(wp, wm1, wm2, g["rows"], grad, lr, b1p, b2p, b1, b2, eps)
    for flags in (0, DET):
        p, m1, m2 = (dev(np.full((h, n), 5.0, np.float32)) for _ in range(3))
        capi.sparse_adam(p, m1, m2, dev(np.array(g["rows"], np.int64)), dev(grad), dev(lr), dev(b1p),
                         dev(b2p), b1, b2, eps, flags=flags)
        # unique rows: both modes are bit-exact against the reference order
        assert_bit_equal(host(p), wp, "G4 param")
        assert_bit_equal(host(m1), wm1, "G4 m1")
        assert_bit_equal(host(m2), wm2, "G4 m2")


@pytest.mark.parametrize("case", SHAPES, ids=lambda c: "n%d_v%d_d%d_%s" % c)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_sparse_adam(capi, case, dtype):
    n, v, d, dist = case
    rng = np.random.default_rng(n * 13 + d)
    rows, grad = make_case(rng, n, v, d, dtype, dist)
    grad *= dtype(0.01)
    p0 = (rng.random((v, d)) * 0.1 - 0.05).astype(dtype)
    m10 = (rng.standard_normal((v, d)) * 1e-3).astype(dtype)
    m20 = (rng.random((v, d)) * 1e-4).astype(dtype)
    lr = np.array([0.001], dtype)
    step = 3
    b1p = np.array([0.9 ** step], dtype)
    b2p = np.array([0.999 ** step], dtype)
    wp, wm1, wm2 = p0.copy(), m10.copy(), m20.copy()
    oracle.adam_sparse(wp, wm1, wm2, rows, grad, lr, b1p, b2p)
    p, m1, m2 = dev(p0), dev(m10), dev(m20)
    capi.sparse_adam(p, m1, m2, dev(rows), dev(grad), dev(lr), dev(b1p), dev(b2p), flags=DET)
    assert_bit_equal(host(p), wp, "adam det param")
    assert_bit_equal(host(m1), wm1, "adam det m1")
    assert_bit_equal(host(m2), wm2, "adam det m2")
    # fast mode: moments follow the merged gradient within the reordering bound; params within
    # 1e-6 relative plus the propagated bound (lr_t/(sqrt(m2)+eps) amplifies tiny m2, so compare
    # against an fp64 evaluation of the same update on the fast merged gradient instead)
    p, m1, m2 = dev(p0), dev(m10), dev(m20)
    capi.sparse_adam(p, m1, m2, dev(rows), dev(grad), dev(lr), dev(b1p), dev(b2p))
    ur, gm, u = capi.merge_add(dev(rows), dev(grad), v)
    u = int(u.item())
    ur, gm = host(ur)[:u], host(gm)[:u]
    ep, em1, em2 = p0.copy(), m10.copy(), m20.copy()
    oracle.adam_sparse(ep, em1, em2, ur, gm, lr, b1p, b2p)   # unique rows -> pure functor
    assert_bit_equal(host(m1), em1, "adam fast m1 == functor(fast merged grad)")
    assert_bit_equal(host(m2), em2, "adam fast m2")
    assert_bit_equal(host(p), ep, "adam fast param")
    # and the fast merged gradient itself is within the bound (checked in test_merge_add);
    # moments are linear in it:
    l1 = np.zeros((v, d))
    _, l1u = oracle.merge_add(rows, np.abs(grad), truth=True)
    l1[ur] = l1u
    rt = FAST_RTOL if dtype == np.float32 else 1e-14
    assert (np.abs(host(m1).astype(np.float64) - wm1) <= rt * (np.abs(wm1) + 0.1 * l1) + 1e-30).all()


# ------------------------------------------------------------------ BASELINE config C2 at full size
def test_config_c2_full_size_step(capi, pkg):
    """1M x 64 table, 4096 x 26 Zipf(1.1) ids, SGD: fwd gather + fused update vs the oracle."""
    v, d, n, a, _ = pkg.synth.CONFIGS["C2_ctr_sgd"]
    ids = pkg.synth.ZipfIds(v, a, seed=0).sample(n)
    assert len(np.unique(ids)) == 25223
    table = pkg.synth.table_values(v, d)
    grad = pkg.synth.grad_values(n, d)
    lr = np.array([0.01], np.float32)
    dt, dids, dgrad, dlr = dev(table), dev(ids), dev(grad), dev(lr)
    out = capi.lookup_table(dt, dids)
    assert_bit_equal(host(out), oracle.lookup_table(table, ids), "C2 gather")
    want = table.copy()
    oracle.sgd_sparse(want, ids, grad, lr)
    pdet = dt.clone()
    capi.sparse_sgd(pdet, dids, dgrad, dlr, flags=DET)
    assert_bit_equal(host(pdet), want, "C2 sgd det")
    pfast = dt.clone()
    capi.sparse_sgd(pfast, dids, dgrad, dlr)
    l1 = np.zeros((v, d))
    ur, l1u = oracle.merge_add(ids, np.abs(grad), truth=True)
    l1[ur] = l1u
    # On the hottest row (13 406 duplicates) the reference's own sequential fp32 chain drifts by
    # more than 1e-6 relative, so the fast mode is held to the fp64 evaluation of the same update
    # and must not be further from it than the reference order is allowed to be.
    truth = table.astype(np.float64)
    truth[ur] -= float(lr[0]) * oracle.merge_add(ids, grad, truth=True)[1]
    bound = FAST_RTOL * (np.abs(table) + 0.01 * l1) + 1e-30
    err_fast = np.abs(host(pfast).astype(np.float64) - truth)
    err_ref = np.abs(want.astype(np.float64) - truth)
    assert (err_fast <= bound).all()
    assert err_fast.max() <= max(err_ref.max(), bound.max())
    # merged index set bit-exact, merged values within bound
    r, val, u = capi.merge_add(dids, dgrad, v)
    u = int(u.item())
    wr, wt = oracle.merge_add(ids, grad, truth=True)
    assert u == 25223 and np.array_equal(host(r)[:u], wr)
    assert (np.abs(host(val)[:u] - wt) <= FAST_RTOL * l1u + 1e-30).all()


# ------------------------------------------------------------------ sparse Adagrad
def test_sparse_adagrad_golden_g5(capi, goldens):
    # test_adagrad_op.py:87-188 — rows [0,4,7,4]: the reference's only duplicate-row MergeAdd test
    import math
    g = goldens["G5_sparse_adagrad"]
    grad = np.ones((4, g["row_numel"]), np.float32)
    for r_, c_, v_ in g["grad_set"]:
        grad[r_, c_] = v_
    for flags in (0, DET):
        p = dev(np.full((g["height"], g["row_numel"]), g["param_fill"], np.float32))
        m = dev(np.full((g["height"], g["row_numel"]), g["moment_fill"], np.float32))
        capi.sparse_adagrad(p, m, dev(np.array(g["rows"], np.int64)), dev(grad), dev(np.array([g["lr"]], np.float32)),
                            g["epsilon"], flags=flags)
        for r_, c_, v_ in g["expect_moment"]:
            assert host(m)[r_, c_] == pytest.approx(v_, abs=1e-6)
        for r_, c_, gr, mm in g["expect_param"]:
            assert host(p)[r_, c_] == pytest.approx(g["param_fill"] - g["lr"] * gr / (math.sqrt(mm) + g["epsilon"]), abs=1e-5)


@pytest.mark.parametrize("case", SHAPES[3:12], ids=lambda c: "n%d_v%d_d%d_%s" % c)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_sparse_adagrad(capi, case, dtype):
    n, v, d, dist = case
    rng = np.random.default_rng(n * 17 + d)
    rows, grad = make_case(rng, n, v, d, dtype, dist)
    grad *= dtype(0.01)
    p0 = (rng.random((v, d)) * 0.1 - 0.05).astype(dtype)
    m0 = (rng.random((v, d)) * 1e-3).astype(dtype)
    lr = np.array([0.01], dtype)
    wp, wm = p0.copy(), m0.copy()
    oracle.adagrad_sparse(wp, wm, rows, grad, lr, 1e-6)
    p, m = dev(p0), dev(m0)
    capi.sparse_adagrad(p, m, dev(rows), dev(grad), dev(lr), 1e-6, flags=DET)
    assert_bit_equal(host(p), wp, "adagrad det param")
    assert_bit_equal(host(m), wm, "adagrad det moment")
    # fast mode == the same functor applied to the fast-mode merged gradient (unique rows -> no reordering left)
    p, m = dev(p0), dev(m0)
    capi.sparse_adagrad(p, m, dev(rows), dev(grad), dev(lr), 1e-6)
    ur, gm, u = capi.merge_add(dev(rows), dev(grad), v)
    u = int(u.item())
    ep, em = p0.copy(), m0.copy()
    oracle.adagrad_sparse(ep, em, host(ur)[:u], host(gm)[:u], lr, 1e-6)
    assert_bit_equal(host(p), ep, "adagrad fast param")
    assert_bit_equal(host(m), em, "adagrad fast moment")


# ------------------------------------------------------------------ bit-exact order mode, long rows
# Rows with more than 64 duplicates leave k_segment_det for k_long_segment_det (one CTA per row and 64-byte column
# slice, 1024-duplicate tiles for 16-byte-aligned widths, 256 otherwise): counts at and around the hand-over and
# the tile sizes, widths with one / several / ragged slices, both load paths, every consumer.
LONG_ROWS = [  # (duplicates of the hot id, other ids, d)
    (64, 5, 4), (65, 5, 4), (255, 0, 7), (256, 3, 7), (257, 3, 31), (1023, 9, 64), (1024, 9, 64), (1025, 0, 64),
    (2048, 50, 16), (3100, 50, 100), (700, 20, 203), (1500, 20, 260),
]


@pytest.mark.parametrize("case", LONG_ROWS, ids=lambda c: "dup%d_other%d_d%d" % c)
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_bit_exact_mode_long_rows(capi, case, dtype):
    dup, other, d = case
    v = 97
    rng = np.random.default_rng(dup * 3 + d)
    rows = np.concatenate([np.full(dup, 41, np.int64), rng.integers(0, v, size=other).astype(np.int64),
                           np.full(dup // 2 + 70, 5, np.int64)])        # a second long row
    rng.shuffle(rows)
    grad = (rng.standard_normal((len(rows), d)) * 0.01).astype(dtype)
    # MergeAdd
    want_rows, want_val = oracle.merge_add(rows, grad)
    r, val, u = capi.merge_add(dev(rows), dev(grad), v, flags=DET)
    u = int(u.item())
    assert u == len(want_rows) and np.array_equal(host(r)[:u], want_rows)
    assert_bit_equal(host(val)[:u], want_val, "merge_add det, long rows")
    # SelectedRowsAddToTensor
    base = rng.standard_normal((v, d)).astype(dtype)
    want = base.copy()
    oracle.selected_rows_add_to_tensor(rows, grad, want)
    t = dev(base)
    capi.selected_rows_add_to_tensor(dev(rows), dev(grad), t, flags=DET)
    assert_bit_equal(host(t), want, "add_to_tensor det, long rows")
    # sgd
    lr = np.array([0.01], dtype)
    want = base.copy()
    oracle.sgd_sparse(want, rows, grad, lr)
    p = dev(base)
    capi.sparse_sgd(p, dev(rows), dev(grad), dev(lr), flags=DET)
    assert_bit_equal(host(p), want, "sgd det, long rows")
    # adam
    m10 = (rng.standard_normal((v, d)) * 1e-3).astype(dtype)
    m20 = (rng.random((v, d)) * 1e-4).astype(dtype)
    b1p, b2p = np.array([0.9 ** 2], dtype), np.array([0.999 ** 2], dtype)
    wp, wm1, wm2 = base.copy(), m10.copy(), m20.copy()
    oracle.adam_sparse(wp, wm1, wm2, rows, grad, lr, b1p, b2p)
    p, m1, m2 = dev(base), dev(m10), dev(m20)
    capi.sparse_adam(p, m1, m2, dev(rows), dev(grad), dev(lr), dev(b1p), dev(b2p), flags=DET)
    assert_bit_equal(host(p), wp, "adam det param, long rows")
    assert_bit_equal(host(m1), wm1, "adam det m1, long rows")
    assert_bit_equal(host(m2), wm2, "adam det m2, long rows")
    # adagrad
    wp, wm = base.copy(), m20.copy()
    oracle.adagrad_sparse(wp, wm, rows, grad, lr, 1e-6)
    p, m = dev(base), dev(m20)
    capi.sparse_adagrad(p, m, dev(rows), dev(grad), dev(lr), 1e-6, flags=DET)
    assert_bit_equal(host(p), wp, "adagrad det param, long rows")
    assert_bit_equal(host(m), wm, "adagrad det moment, long rows")
    # one plan, several consumers on the same workspace (the fused wrappers' pattern): the long-row list and its
    # counters live in scratch of the fast path and must be left the way the fast path expects them
    ws = capi.alloc_workspace(len(rows), d, grad.itemsize)
    drows, dgrad = dev(rows), dev(grad)
    capi.plan_rows(drows, v, ws)
    want = base.copy()
    oracle.sgd_sparse(want, rows, grad, lr)
    oracle.sgd_sparse(want, rows, grad, lr)
    p = dev(base)
    capi.sparse_sgd(p, drows, dgrad, dev(lr), workspace=ws, flags=DET | capi.PLAN_READY)
    capi.sparse_sgd(p, drows, dgrad, dev(lr), workspace=ws, flags=DET | capi.PLAN_READY)
    assert_bit_equal(host(p), want, "sgd det twice on one plan, long rows")
    pf, pg = dev(base), dev(base)
    capi.sparse_sgd(pf, drows, dgrad, dev(lr), workspace=ws, flags=capi.PLAN_READY)      # fast, after det, same plan
    capi.sparse_sgd(pg, drows, dgrad, dev(lr))                                            # fast, fresh workspace
    assert_bit_equal(host(pf), host(pg), "fast mode after bit-exact mode on one plan")


def test_bit_exact_mode_run_of_exactly_the_handover_length(capi):
    for n in (63, 64, 65):
        rows = np.full(n, 3, np.int64)
        grad = np.random.default_rng(n).standard_normal((n, 12)).astype(np.float32)
        base = np.zeros((5, 12), np.float32)
        want = base.copy()
        lr = np.array([0.5], np.float32)
        oracle.sgd_sparse(want, rows, grad, lr)
        p = dev(base)
        capi.sparse_sgd(p, dev(rows), dev(grad), dev(lr), flags=DET)
        assert_bit_equal(host(p), want, "sgd det, %d duplicates and nothing else" % n)


# ------------------------------------------------------------------ committed outputs of the reference kernels
def test_bit_exact_mode_matches_committed_outputs_of_the_reference_kernels(capi, ref_fixtures):
    """The CUDA path against tests/golden/ref_outputs.npz directly — what the reference's unmodified CPU kernels
    produced on these seeded inputs — with no restatement in between."""
    mod, store = ref_fixtures
    for case in mod.CASES:
        for dt in mod.DTYPES:
            x = mod.inputs(*case, dt)
            v = x["table"].shape[0]
            want = lambda name: store[mod.key(case, dt, name)]
            rows, value, lr = dev(x["rows"]), dev(x["value"]), dev(x["lr"])
            assert np.array_equal(mod.bits(host(capi.lookup_table(dev(x["table"]), rows))), want("lookup"))
            out = capi.lookup_table(dev(x["table"]), rows, int(x["rows"][-1]))
            assert np.array_equal(mod.bits(host(out)), want("lookup_pad"))
            r, val, u = capi.merge_add(rows, value, v, flags=DET)
            u = int(u.item())
            assert np.array_equal(host(r)[:u], want("merge_rows"))
            assert np.array_equal(mod.bits(host(val)[:u]), want("merge_value"))
            t = dev(x["table"])
            capi.selected_rows_add_to_tensor(rows, value, t, flags=DET)
            assert np.array_equal(mod.bits(host(t)), want("add_to_tensor"))
            p = dev(x["table"])
            capi.sparse_sgd(p, rows, value, lr, flags=DET)
            assert np.array_equal(mod.bits(host(p)), want("sgd"))
            p, m1, m2 = dev(x["table"]), dev(x["m1"]), dev(x["m2"])
            capi.sparse_adam(p, m1, m2, rows, value, lr, dev(x["b1p"]), dev(x["b2p"]), flags=DET)
            assert np.array_equal(mod.bits(host(p)), want("adam_p"))
            assert np.array_equal(mod.bits(host(m1)), want("adam_m1"))
            assert np.array_equal(mod.bits(host(m2)), want("adam_m2"))


def test_cluster_sort_opt_in_path_matches_default():
    """B200EMB_CLUSTER_SORT=1 routes small plans through the single 16-CTA-cluster DSMEM sort
    (csrc/cluster_sort.cu).  The env var is read once per process, so the check runs in a child process:
    merged rows and deterministic-mode SGD must equal the oracle bit for bit there too."""
    import os
    import subprocess
    import sys
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import oracle, __graft_entry__ as ge
capi = ge.load_package().capi
rng = np.random.default_rng(3)
for n, v, d in [(7, 50, 8), (5000, 4096, 32), (106496, 1000000, 64), (139000, 1 << 27, 16)]:
    w = np.arange(1, min(v, 200000) + 1, dtype=np.float64) ** -1.1
    cdf = np.cumsum(w) / w.sum()
    rows = ((np.searchsorted(cdf, rng.random(n)) * 2654435761 + 12345) %% v).astype(np.int64)
    grad = (rng.standard_normal((n, d)) * 0.01).astype(np.float32)
    wr, wv = oracle.merge_add(rows, grad)
    r, val, u = capi.merge_add(torch.from_numpy(rows).cuda(), torch.from_numpy(grad).cuda(), v, flags=1)
    u = int(u.item())
    assert u == len(wr) and np.array_equal(r.cpu().numpy()[:u], wr)
    assert np.array_equal(val.cpu().numpy()[:u].view(np.uint32), wv.view(np.uint32))
print("cluster-sort-ok")
''' % root
    env = dict(os.environ, B200EMB_CLUSTER_SORT="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "cluster-sort-ok" in out.stdout, out.stderr[-2000:]


def test_tma_bulk_gather_opt_in_path_is_bit_exact():
    """B200EMB_GATHER_BULK=1 routes rows of >= 512 B through the TMA bulk-copy gather (cp.async.bulk +
    mbarrier, csrc/gather.cu k_gather_bulk).  Checked in a child process (env var is read once)."""
    import os
    import subprocess
    import sys
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import oracle, __graft_entry__ as ge
capi = ge.load_package().capi
rng = np.random.default_rng(5)
for d, dtype in [(128, np.float32), (256, np.float32), (1028, np.float32), (64, np.float64), (3000, np.float32)]:
    v, n = 2000, 70001
    table = rng.standard_normal((v, d)).astype(dtype)
    ids = rng.integers(0, v, size=n).astype(np.int64)
    ids[5] = v + 3
    pad = int(ids[11])
    status = capi.new_status()
    got = capi.lookup_table(torch.from_numpy(table).cuda(), torch.from_numpy(ids).cuda(), pad, status=status).cpu().numpy()
    want_ids = ids.copy(); want_ids[5] = pad
    want = oracle.lookup_table(table, want_ids, pad)
    assert np.array_equal(got, want), d
    assert int(status.item()) & 1
print("bulk-gather-ok")
''' % root
    env = dict(os.environ, B200EMB_GATHER_BULK="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "bulk-gather-ok" in out.stdout, out.stderr[-2000:]
