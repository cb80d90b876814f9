"""Pins the CPU oracle against the reference's own known-answer tests (plus G11).
CPU only.  Each test names the reference test it restates."""
import math

import numpy as np
import pytest

import oracle


def _filled(rows, numel, fill, sets=()):
    a = np.full((len(rows), numel), fill, dtype=np.float32)
    for r, c, v in sets:
        a[r, c] = v
    return a


def test_g1_g2_selected_rows_add(goldens):
    # operators/math/selected_rows_functor_test.cc:20-221
    g = goldens["G1_G2_selected_rows_add"]
    n = g["row_numel"]
    v1 = _filled(g["rows1"], n, g["fill1"])
    v2 = _filled(g["rows2"], n, g["fill2"])
    rows, value = oracle.sum_selected_rows([(g["rows1"], v1), (g["rows2"], v2)])
    assert rows.tolist() == g["expect_rows"]
    for r, c, v in g["expect_concat"]:
        assert value[r, c] == v
    t = np.full((g["height"], n), g["tensor_fill"], dtype=np.float32)
    dense = oracle.selected_rows_add_tensor(rows, value, t)          # cpu_add
    t2 = t.copy()
    oracle.selected_rows_add_to_tensor(rows, value, t2)              # cpu_add_to
    for r, c, v in g["expect_dense"]:
        assert dense[r, c] == v
        assert t2[r, c] == v
    assert np.array_equal(dense, t2)


def test_g3_sparse_sgd(goldens):
    # test_sgd_op.py:38-99
    g = goldens["G3_sparse_sgd"]
    grad = _filled(g["rows"], g["row_numel"], g["grad_fill"], g["grad_set"])
    param = np.full((g["height"], g["row_numel"]), g["param_fill"], dtype=np.float32)
    oracle.sgd_sparse(param, g["rows"], grad, [g["lr"]])
    for r, c, v in g["expect"]:
        assert param[r, c] == pytest.approx(v, abs=1e-7)


def _adam_step_sparse_numpy(g, grad):
    # numpy model of test_adam_op.py:197-229 (float64 like the reference test)
    h, n = g["height"], g["row_numel"]
    b1, b2, eps = g["beta1"], g["beta2"], g["epsilon"]
    p = np.full((h, n), g["param_fill"], np.float32)
    m1 = np.full((h, n), g["moment1_fill"], np.float32)
    m2 = np.full((h, n), g["moment2_fill"], np.float32)
    b1p = np.array([b1 ** g["beta_pow_exponent"]], np.float32)
    b2p = np.array([b2 ** g["beta_pow_exponent"]], np.float32)
    lr = np.full((1,), g["lr"], np.float32)
    po, m1o, m2o = np.zeros((h, n)), np.zeros((h, n)), np.zeros((h, n))
    for idx, r in enumerate(g["rows"]):
        m1o[r] = b1 * m1[r] + (1 - b1) * grad[idx]
        m2o[r] = b2 * m2[r] + (1 - b2) * np.square(grad[idx])
        lr_t = lr * np.sqrt(1 - b2p) / (1 - b1p)
        po[r] = p[r] - lr_t * (m1o[r] / (np.sqrt(m2o[r]) + eps))
    return (p, m1, m2, lr, b1p, b2p), (po, m1o, m2o)


@pytest.mark.parametrize("reference_form", [True, False])
def test_g4_sparse_adam(goldens, reference_form):
    # test_adam_op.py:232-313, rel 1e-5 on touched rows
    g = goldens["G4_sparse_adam"]
    grad = _filled(g["rows"], g["row_numel"], g["grad_fill"], g["grad_set"])
    (p, m1, m2, lr, b1p, b2p), expect = _adam_step_sparse_numpy(g, grad)
    oracle.adam_sparse(p, m1, m2, g["rows"], grad, lr, b1p, b2p, g["beta1"], g["beta2"],
                       g["epsilon"], reference_form=reference_form)
    for got, want in zip((p, m1, m2), expect):
        for r in g["rows"]:
            rel = np.abs((got[r] - want[r]) / got[r])
            assert rel.max() < g["rel_tol"]
    # untouched rows keep their values ("lazy" adam): adam_op.h:199-219 only writes rows_[i]
    untouched = [r for r in range(g["height"]) if r not in g["rows"]]
    assert np.all(p[untouched] == g["param_fill"])
    assert np.all(m1[untouched] == g["moment1_fill"])


def test_g5_sparse_adagrad_duplicate_rows(goldens):
    # test_adagrad_op.py:87-188 — rows [0,4,7,4]
    g = goldens["G5_sparse_adagrad"]
    grad = _filled(g["rows"], g["row_numel"], g["grad_fill"], g["grad_set"])
    param = np.full((g["height"], g["row_numel"]), g["param_fill"], np.float32)
    moment = np.full((g["height"], g["row_numel"]), g["moment_fill"], np.float32)
    u = oracle.adagrad_sparse(param, moment, g["rows"], grad, [g["lr"]], g["epsilon"])
    assert u == 3
    for r, c, v in g["expect_moment"]:
        assert moment[r, c] == pytest.approx(v, abs=1e-6)
    for r, c, gr, m in g["expect_param"]:
        want = g["param_fill"] - g["lr"] * gr / (math.sqrt(m) + g["epsilon"])
        assert param[r, c] == pytest.approx(want, abs=1e-5)
    # MergeAdd of the duplicate: rows sorted unique, row 4 = 1+1
    mr, mv = oracle.merge_add(g["rows"], grad, reference_form=True)
    assert mr.tolist() == [0, 4, 7]
    assert np.all(mv[1] == 2.0)


@pytest.mark.parametrize("shape_idx", [0, 1])
@pytest.mark.parametrize("padding", [None, "pos", "neg"])
def test_g6_lookup_table(goldens, shape_idx, padding):
    # test_lookup_table_op.py:25-83
    g = goldens["G6_lookup_table"]
    rng = np.random.default_rng(shape_idx)
    table = rng.random(g["table_shape"]).astype(np.float32)
    ids = rng.integers(0, g["table_shape"][0], size=g["ids_shapes"][shape_idx]).astype(np.int64)
    flat = ids.reshape(-1)
    expect = table[flat]
    pad = -1
    if padding is not None:
        pad = int(flat[rng.integers(0, flat.size)])
        expect = expect.copy()
        expect[flat == pad] = 0
        # the python layer maps a negative padding_idx to size[0]+padding_idx (layers/nn.py:276-277)
        if padding == "neg":
            neg = pad - g["table_shape"][0]
            pad = g["table_shape"][0] + neg
    out = oracle.lookup_table(table, ids, pad)
    assert np.array_equal(out, expect)


def test_lookup_table_rejects_out_of_range():
    # lookup_table_op.h:59-60
    table = np.zeros((5, 3), np.float32)
    with pytest.raises(IndexError):
        oracle.lookup_table(table, [0, 5])
    with pytest.raises(IndexError):
        oracle.lookup_table(table, [-2])


def test_lookup_table_selected_rows_table():
    # test_lookup_table_op.py:86-150 (W as SelectedRows, CPU only)
    rows = [0, 1, 2, 3, 4, 5, 6]
    w = np.ones((7, 12), np.float32) * np.arange(7, dtype=np.float32)[:, None]
    ids = np.array([0, 4, 3, 5], np.int64)
    out = oracle.lookup_table_selected_rows(rows, w, ids)
    for k, i in enumerate(ids):
        assert np.all(out[k] == i)


def test_g7_sum_selected_rows_with_empties(goldens):
    # test_sum_op.py:47-105
    g = goldens["G7_sum_selected_rows"]
    n = g["row_numel"]
    full = np.ones((len(g["rows"]), n), np.float32) * np.arange(len(g["rows"]), dtype=np.float32)[:, None]
    empty = np.ones((0, n), np.float32)
    for case in g["cases"]:
        ins = [(g["rows"], full) if has else ([], empty) for has in case]
        rows, value = oracle.sum_selected_rows(ins)
        assert len(rows) == 7 * sum(case)
        assert value.shape == (7 * sum(case), n)
        if sum(case):
            assert rows.tolist() == g["rows"] * sum(case)
            assert np.array_equal(value, np.concatenate([full] * sum(case)))


@pytest.mark.parametrize("ndev", [2, 8])
def test_g8_gather_selected_rows(goldens, ndev):
    # framework/details/reduce_op_handle_test.cc:153-209 (8 fake CPU "devices" at :74-79)
    g = goldens["G8_gather_selected_rows"]
    rows = g["rows"]
    value = np.arange(len(rows) * g["width"], dtype=np.float32).reshape(len(rows), g["width"])
    out_rows, out_value = oracle.gather_selected_rows([(rows, value)] * ndev)
    assert len(out_rows) == ndev * len(rows)
    for k in range(len(out_rows)):
        assert out_rows[k] == rows[k % len(rows)]
    flat = out_value.reshape(-1)
    assert np.array_equal(flat, np.tile(value.reshape(-1), ndev))


def test_g9_split_merge_ids(goldens):
    g = goldens["G9_split_merge_ids"]
    shards = oracle.split_ids(g["ids"], g["shard_num"])
    assert [s.tolist() for s in shards] == g["expect_shards"]
    # SelectedRows flavour: test_split_ids_op.py:41-78
    rows = g["sr_rows"]
    value = np.array([[r + j for j in range(g["sr_row_numel"])] for r in rows], np.float32)
    outs = oracle.split_ids_selected_rows(rows, value, g["shard_num"])
    assert [o[0].tolist() for o in outs] == g["expect_sr_rows"]
    for orows, ovals in outs:
        for r, v in zip(orows, ovals):
            assert v.tolist() == [r + j for j in range(g["sr_row_numel"])]
    # merge_ids: test_merge_ids_op.py:22-34
    xs = [np.array(x, np.float32).reshape(-1, 2) for x in g["merge_x"]]
    out = oracle.merge_ids(g["ids"], xs)
    assert np.array_equal(out, np.array(g["merge_expect"], np.float32))
    # split_selected_rows: test_split_selected_rows_op.py:36-83
    v = _filled(g["ssr_rows"], g["ssr_row_numel"], 1.0, g["ssr_value_set"])
    outs = oracle.split_selected_rows(g["ssr_rows"], v, g["ssr_height_sections"])
    for k, want in g["ssr_expect_rows"].items():
        assert outs[int(k)][0].tolist() == want
    for sec, r, c, val in g["ssr_expect_value"]:
        assert outs[sec][1][r, c] == val


def test_merge_add_forms_are_bit_identical():
    # the O(N log N) form used at BASELINE sizes must equal the reference's O(N*U) form
    rng = np.random.default_rng(7)
    for n, v, d in [(1, 3, 4), (257, 19, 8), (4000, 300, 33)]:
        rows = rng.integers(0, v, size=n)
        value = rng.standard_normal((n, d)).astype(np.float32)
        r1, v1 = oracle.merge_add(rows, value, reference_form=True)
        r2, v2 = oracle.merge_add(rows, value, reference_form=False)
        assert np.array_equal(r1, r2) and np.array_equal(v1.view(np.uint32), v2.view(np.uint32))
        assert np.array_equal(r1, np.unique(rows))
        rt, vt = oracle.merge_add(rows, value, truth=True)
        assert np.allclose(v1, vt, rtol=0, atol=1e-4)


def test_merge_add_negative_zero_segment_gives_positive_zero():
    # accumulation starts from +0.0
    rows = [3, 3]
    value = np.full((2, 2), -0.0, np.float32)
    _, mv = oracle.merge_add(rows, value, reference_form=True)
    assert not np.signbit(mv).any()


def test_sgd_with_duplicates_is_sequential_not_merged():
    # sgd_op.h:74-80 applies p -= lr*g once per occurrence (two roundings each)
    lr = np.float32(0.1)
    g = np.array([[1.0000001], [3.0000002], [7.0000005]], np.float32)
    p = np.array([[0.3]], np.float32)
    want = p.copy()
    for i in range(3):
        want[0, 0] = np.float32(want[0, 0] - np.float32(lr * g[i, 0]))
    oracle.sgd_sparse(p, [0, 0, 0], g, [lr])
    assert p[0, 0] == want[0, 0]


def test_empty_grad_is_a_no_op():
    p = np.ones((4, 3), np.float32)
    oracle.sgd_sparse(p, [], np.zeros((0, 3), np.float32), [0.5])
    assert np.all(p == 1)
    m1, m2 = p.copy(), p.copy()
    assert oracle.adam_sparse(p, m1, m2, [], np.zeros((0, 3), np.float32), [0.5], [0.9], [0.99]) == 0
    assert np.all(p == 1) and np.all(m1 == 1)


def test_lookup_table_grad_sparse_keeps_duplicates_and_order():
    # lookup_table_op.h:114-119 / SURVEY Appendix A-1
    ids = np.array([[5], [2], [5], [0]], np.int64)
    dout = np.arange(8, dtype=np.float32).reshape(4, 2)
    rows, value = oracle.lookup_table_grad_sparse(ids, dout)
    assert rows.tolist() == [5, 2, 5, 0]
    assert np.array_equal(value, dout)
    dense = oracle.lookup_table_grad_dense(ids, dout, 6)
    assert dense[5].tolist() == [0 + 4, 1 + 5]


def test_g11_regulariser_ops(goldens):
    g = goldens["G11_regulariser_ops"]
    # test_extract_rows_op.py:28-53
    assert [r[0] for r in oracle.extract_rows(g["extract_rows"]).tolist()] == g["extract_rows"]
    rng = np.random.default_rng(11)
    # test_scale_op.py:38-78 — scale on a SelectedRows: with a table whose row r is in_array[i] the chain
    # lookup_table -> scale must give in_array * scale exactly
    rows = g["scale_rows"]
    table = np.zeros((g["scale_height"], g["scale_row_numel"]), np.float32)
    in_array = rng.random((len(rows), g["scale_row_numel"])).astype(np.float32)
    table[rows] = in_array
    assert (oracle.sparse_decay(table, rows, g["scale"]) == in_array * np.float32(g["scale"])).all()
    # test_scale_op.py:26-31 (dense, scale -2.3) and test_sign_op.py:23-30 through the same chain
    x = rng.random(tuple(g["dense_shape"])).astype(np.float32)
    ident = np.arange(x.shape[0])
    assert (oracle.sparse_decay(x, ident, g["dense_scale"]) == x * np.float32(g["dense_scale"])).all()
    lo, hi = g["sign_range"]
    y = rng.uniform(lo, hi, tuple(g["sign_shape"])).astype(np.float32)
    y[0, :3] = [0.0, -0.0, np.nan]
    want = np.sign(y)
    want[0, 2] = 0.0           # Eigen's (a>0)-(a<0) gives 0 for NaN where numpy propagates it
    assert (oracle.sparse_decay(y, ident, 1.0, mode=1) == want).all()
    # the whole chain with the final elementwise_add, duplicates each getting their own decay term
    grad = rng.standard_normal((4, 12)).astype(np.float32)
    tab = rng.standard_normal((8, 12)).astype(np.float32)
    r = np.array(g["extract_rows"])
    got = oracle.sparse_decay(tab, r, 0.01, 0, addend=grad)
    assert (got == grad + np.float32(0.01) * tab[r]).all()
    with pytest.raises(IndexError):
        oracle.sparse_decay(tab, [8], 0.01)


def test_oracle_matches_committed_outputs_of_the_reference_kernels(ref_fixtures):
    """tests/golden/ref_outputs.npz holds what the reference's unmodified CPU kernels (oracle/_ref) produced on
    seeded inputs; the restatement must reproduce every bit — on any machine, with or without /root/reference."""
    mod, store = ref_fixtures
    checked = 0
    for case in mod.CASES:
        for dt in mod.DTYPES:
            x = mod.inputs(*case, dt)
            got = mod.outputs(oracle, x, adam_kwargs={"reference_form": True})
            for name, arr in got.items():
                want = store[mod.key(case, dt, name)]
                assert np.array_equal(mod.bits(arr), want), (case, np.dtype(dt).name, name)
                checked += 1
            # the restatement's default Adam form and both MergeAdd forms give the same bits as well
            a = [x["table"].copy(), x["m1"].copy(), x["m2"].copy()]
            oracle.adam_sparse(a[0], a[1], a[2], x["rows"], x["value"], x["lr"], x["b1p"], x["b2p"])
            assert np.array_equal(mod.bits(a[0]), store[mod.key(case, dt, "adam_p")])
            _, mv = oracle.merge_add(x["rows"], x["value"], reference_form=True)
            assert np.array_equal(mod.bits(mv), store[mod.key(case, dt, "merge_value")])
    assert checked == len(store.files)
