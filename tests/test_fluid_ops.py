"""The reference's own Python op tests for this path, restated against the B200 OpKernel wrappers
(paddle-1_b200/fluid/*.cu, registered with REGISTER_OP_CUDA_KERNEL and dispatched like
OperatorWithKernel::RunImpl).  `core` / `Operator` below are paddle-1_b200/fluid_lite.py, a mirror of the
names the reference tests import.  Source tests (under python/paddle/fluid/tests/unittests/):
test_lookup_table_op.py:25-83, test_sgd_op.py:38-99, test_adam_op.py:197-313, test_sum_op.py:47-105."""
import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def core(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import importlib
    mod = importlib.import_module("paddle1_b200.fluid_lite")
    mod.lib()
    return mod


def test_kernels_are_registered(core):
    for op in ("lookup_table", "lookup_table_grad", "sgd", "adam", "adagrad"):
        assert core.registered_kernels(op) == 2        # float + double, CUDAPlace
    assert core.registered_kernels("sum") == 4         # + int, int64 (operators/sum_op.cu:16-20)


def test_no_cpu_kernel(core):
    scope = core.Scope()
    scope.var("W").get_tensor().set(np.ones((4, 3), np.float32), core.CPUPlace())
    scope.var("Ids").get_tensor().set(np.zeros((2, 1), np.int64), core.CPUPlace())
    with pytest.raises(core.FluidLiteError):
        core.Operator("lookup_table", W="W", Ids="Ids", Out="Out").run(scope, core.CPUPlace())


@pytest.mark.parametrize("ids_shape", [(4, 1), (2, 4, 5, 1)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_lookup_table_op(core, ids_shape, dtype):
    # TestLookupTableOp / TestLookupTableOpWithTensorIds / ...WithPadding (test_lookup_table_op.py:25-83)
    place = core.CUDAPlace(0)
    table = np.random.random((17, 31)).astype(dtype)
    ids = np.random.randint(0, 17, ids_shape).astype("int64")
    for padding_idx in (-1, int(np.random.choice(ids.flatten()))):
        scope = core.Scope()
        scope.var("W").get_tensor().set(table, place)
        scope.var("Ids").get_tensor().set(ids, place)
        core.Operator("lookup_table", W="W", Ids="Ids", Out="Out", padding_idx=padding_idx).run(scope, place)
        out = np.array(scope.var("Out").get_tensor())
        expect = table[ids.flatten()].reshape(ids.shape[:-1] + (31,)).copy()
        if padding_idx != -1:
            expect[np.squeeze(ids == padding_idx)] = np.zeros(31)
        assert out.shape == expect.shape and np.array_equal(out, expect)


def test_lookup_table_grad_sparse_then_sum_then_sgd_word2vec_shape(core):
    # the op sequence of an embedding(is_sparse=True) step: 4 lookups on a shared table,
    # 4 lookup_table_grad -> sum (SelectedRows) -> sgd; config C1 shapes (dict 2073, dim 32, batch 100)
    place = core.CUDAPlace(0)
    V, D, B = 2073, 32, 100
    rng = np.random.default_rng(0)
    table = (rng.random((V, D)) * 0.1 - 0.05).astype(np.float32)
    ids = [rng.integers(0, V, (B, 1)).astype(np.int64) for _ in range(4)]
    douts = [(rng.standard_normal((B, D)) * 0.01).astype(np.float32) for _ in range(4)]
    scope = core.Scope()
    scope.var("shared_w").get_tensor().set(table, place)
    scope.var("lr").get_tensor().set(np.array([0.001], np.float32), place)
    for k in range(4):
        scope.var("ids%d" % k).get_tensor().set(ids[k], place)
        core.Operator("lookup_table", W="shared_w", Ids="ids%d" % k, Out="emb%d" % k, is_sparse=True).run(scope, place)
        assert np.array_equal(np.array(scope.var("emb%d" % k).get_tensor()), table[ids[k].flatten()])
        scope.var("emb%d@GRAD" % k).get_tensor().set(douts[k], place)
        scope.var("g%d" % k).get_selected_rows()
        op = core.Operator("lookup_table_grad", W="shared_w", Ids="ids%d" % k, is_sparse=True,
                           **{"Out@GRAD": "emb%d@GRAD" % k, "W@GRAD": "g%d" % k})
        op.run(scope, place)
        g = scope.var("g%d" % k).get_selected_rows()
        assert g.rows() == ids[k].flatten().tolist() and g.height() == V      # duplicates kept, input order
        assert np.array_equal(np.array(g.get_tensor()), douts[k])
    scope.var("shared_w@GRAD").get_selected_rows().set_height(V)
    core.Operator("sum", X=["g0", "g1", "g2", "g3"], Out="shared_w@GRAD").run(scope, place)
    total = scope.var("shared_w@GRAD").get_selected_rows()
    all_ids = np.concatenate([i.flatten() for i in ids])
    all_g = np.concatenate(douts)
    assert total.rows() == all_ids.tolist() and np.array_equal(np.array(total.get_tensor()), all_g)
    core.Operator("sgd", Param="shared_w", Grad="shared_w@GRAD", ParamOut="shared_w", LearningRate="lr").run(scope, place)
    got = np.array(scope.var("shared_w").get_tensor())
    truth = table.astype(np.float64)
    rows, merged = oracle.merge_add(all_ids, all_g, truth=True)
    truth[rows] -= 0.001 * merged
    assert np.abs(got - truth).max() < 1e-7
    untouched = np.setdiff1d(np.arange(V), all_ids)
    assert np.array_equal(got[untouched], table[untouched])


def _embedding_step(core, fused, lookups, V, D, B, optimizer, seed=0):
    """lookup_table x lookups -> lookup_table_grad x lookups -> [sum] -> sgd|adam through Operator.run; returns
    (forward outputs, updated table)."""
    place = core.CUDAPlace(0)
    rng = np.random.default_rng(seed)
    table = (rng.random((V, D)) * 0.1 - 0.05).astype(np.float32)
    ids = [(rng.zipf(1.3, (B, 1)) % V).astype(np.int64) for _ in range(lookups)]
    douts = [(rng.standard_normal((B, D)) * 0.01).astype(np.float32) for _ in range(lookups)]
    core.set_fused(fused)
    try:
        scope = core.Scope()
        scope.var("w").get_tensor().set(table, place)
        scope.var("lr").get_tensor().set(np.array([0.01], np.float32), place)
        outs = []
        for k in range(lookups):
            scope.var("ids%d" % k).get_tensor().set(ids[k], place)
            core.Operator("lookup_table", W="w", Ids="ids%d" % k, Out="emb%d" % k, is_sparse=True).run(scope, place)
            outs.append(np.array(scope.var("emb%d" % k).get_tensor()))
        for k in range(lookups):
            scope.var("emb%d@GRAD" % k).get_tensor().set(douts[k], place)
            scope.var("g%d" % k).get_selected_rows()
            core.Operator("lookup_table_grad", W="w", Ids="ids%d" % k, is_sparse=True,
                          **{"Out@GRAD": "emb%d@GRAD" % k, "W@GRAD": "g%d" % k}).run(scope, place)
        grad = "g0"
        if lookups > 1:
            scope.var("w@GRAD").get_selected_rows().set_height(V)
            core.Operator("sum", X=["g%d" % k for k in range(lookups)], Out="w@GRAD").run(scope, place)
            grad = "w@GRAD"
        if optimizer == "sgd":
            core.Operator("sgd", Param="w", Grad=grad, ParamOut="w", LearningRate="lr").run(scope, place)
        else:
            for name, val in (("m1", np.zeros((V, D), np.float32)), ("m2", np.zeros((V, D), np.float32)),
                              ("b1p", np.array([0.9], np.float32)), ("b2p", np.array([0.999], np.float32))):
                scope.var(name).get_tensor().set(val, place)
            core.Operator("adam", Param="w", Grad=grad, LearningRate="lr", Moment1="m1", Moment2="m2", Beta1Pow="b1p",
                          Beta2Pow="b2p", ParamOut="w", Moment1Out="m1", Moment2Out="m2").run(scope, place)
        return outs, np.array(scope.var("w").get_tensor()), table, ids, douts
    finally:
        core.set_fused(False)


@pytest.mark.parametrize("lookups,optimizer,B", [(1, "sgd", 40000), (4, "sgd", 3000), (1, "adam", 20000)])
def test_fused_mode_equals_materialised_mode(core, lookups, optimizer, B):
    """FLAGS_b200emb_fused (b200_side_table.h): the gradient stays on the device — no rows() round trip, no
    copies, sort plan started by the forward op — and the step gives the same bits as the materialised mode
    (same plan, same reduction tree), which in turn is checked against the oracle."""
    V, D = 50021, 64
    outs_a, w_a, table, ids, douts = _embedding_step(core, False, lookups, V, D, B, optimizer)
    outs_b, w_b, _, _, _ = _embedding_step(core, True, lookups, V, D, B, optimizer)
    for a, b in zip(outs_a, outs_b):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.array_equal(w_a.view(np.uint32), w_b.view(np.uint32))
    if optimizer == "sgd":
        all_ids = np.concatenate([i.flatten() for i in ids])
        all_g = np.concatenate(douts)
        rows, merged = oracle.merge_add(all_ids, all_g, truth=True)
        _, l1 = oracle.merge_add(all_ids, np.abs(all_g), truth=True)
        truth = table.astype(np.float64)
        truth[rows] -= 0.01 * merged
        bound = 1e-6 * np.abs(table).astype(np.float64)
        bound[rows] += 1e-6 * 0.01 * l1
        assert (np.abs(w_b - truth) <= bound + 1e-30).all()


def test_lookup_table_grad_dense_branch(core):
    """is_sparse=False (lookup_table_op.h:135-157): W@GRAD is a dense [V, D] tensor, zero except the looked-up
    rows, duplicates accumulated."""
    place = core.CUDAPlace(0)
    rng = np.random.default_rng(3)
    V, D, B = 301, 24, 500
    table = rng.random((V, D)).astype(np.float32)
    ids = rng.integers(0, V // 3, (B, 1)).astype(np.int64)
    dout = (rng.standard_normal((B, D)) * 0.1).astype(np.float32)
    scope = core.Scope()
    scope.var("W").get_tensor().set(table, place)
    scope.var("Ids").get_tensor().set(ids, place)
    scope.var("Out@GRAD").get_tensor().set(dout, place)
    core.Operator("lookup_table_grad", W="W", Ids="Ids", is_sparse=False,
                  **{"Out@GRAD": "Out@GRAD", "W@GRAD": "W@GRAD"}).run(scope, place)
    got = np.array(scope.var("W@GRAD").get_tensor())
    want = oracle.lookup_table_grad_dense(ids.flatten(), dout, V)
    assert got.shape == (V, D)
    truth = np.zeros((V, D))
    np.add.at(truth, ids.flatten(), dout.astype(np.float64))
    l1 = np.zeros((V, D))
    np.add.at(l1, ids.flatten(), np.abs(dout).astype(np.float64))
    assert (np.abs(got - truth) <= 1e-6 * l1 + 1e-30).all()
    assert np.allclose(got, want, rtol=1e-5, atol=1e-7)
    assert not got[V // 3:].any()


@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_sum_op_integer_tensors(core, dtype):
    """sum is registered for int and int64 too (operators/sum_op.cu:16-20): dense LoDTensor inputs."""
    place = core.CUDAPlace(0)
    rng = np.random.default_rng(4)
    xs = [rng.integers(-1000, 1000, (37, 5)).astype(dtype) for _ in range(3)]
    scope = core.Scope()
    for k, x in enumerate(xs):
        scope.var("x%d" % k).get_tensor().set(x, place)
    scope.var("out").get_tensor().set(np.zeros((37, 5), dtype), place)
    core.Operator("sum", X=["x0", "x1", "x2"], Out="out").run(scope, place)
    assert np.array_equal(np.array(scope.var("out").get_tensor()), xs[0] + xs[1] + xs[2])


def test_sparse_sgd_op(core):
    # TestSparseSGDOp.check_with_place (test_sgd_op.py:38-99), verbatim values
    place = core.CUDAPlace(0)
    scope = core.Scope()
    height, rows, row_numel = 10, [0, 4, 7], 12
    grad_selected_rows = scope.var('Grad').get_selected_rows()
    grad_selected_rows.set_height(height)
    grad_selected_rows.set_rows(rows)
    np_array = np.ones((len(rows), row_numel)).astype("float32")
    np_array[0, 0] = 2.0
    np_array[2, 8] = 4.0
    grad_selected_rows.get_tensor().set(np_array, place)
    scope.var('Param').get_tensor().set(np.full((height, row_numel), 5.0).astype("float32"), place)
    scope.var('LearningRate').get_tensor().set(np.full((1), 2.0).astype("float32"), place)
    sgd_op = core.Operator("sgd", Param='Param', Grad='Grad', ParamOut='Param', LearningRate='LearningRate')
    sgd_op.run(scope, place)
    result_array = np.array(scope.var('Param').get_tensor())
    assert result_array[rows[0], 0] == pytest.approx(1.0)     # 5.0 - 2.0 * 2.0
    assert result_array[rows[0], 2] == pytest.approx(3.0)     # 5.0 - 2.0 * 1.0
    assert result_array[1, 0] == pytest.approx(5.0)
    assert result_array[rows[1], 10] == pytest.approx(3.0)
    assert result_array[5, 8] == pytest.approx(5.0)
    assert result_array[rows[2], 1] == pytest.approx(3.0)
    assert result_array[rows[2], 8] == pytest.approx(-3.0)    # 5.0 - 2.0 * 4.0


def test_sparse_adam_op(core):
    # TestSparseAdamOp (test_adam_op.py:232-313) with the numpy model adam_step_sparse (:197-229)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    beta1, beta2, epsilon = 0.78, 0.836, 1e-4
    height, rows, row_numel = 10, [0, 4, 7], 12
    dense_inputs = {
        "Param": np.full((height, row_numel), 5.0).astype("float32"),
        "Moment1": np.full((height, row_numel), 5.0).astype("float32"),
        "Moment2": np.full((height, row_numel), 5.0).astype("float32"),
        'Beta1Pow': np.array([beta1 ** 10]).astype("float32"),
        'Beta2Pow': np.array([beta2 ** 10]).astype("float32"),
        "LearningRate": np.full((1), 2.0).astype("float32"),
    }
    g = scope.var('Grad').get_selected_rows()
    g.set_height(height)
    g.set_rows(rows)
    np_array = np.ones((len(rows), row_numel)).astype("float32")
    np_array[0, 0] = 2.0
    np_array[2, 8] = 4.0
    g.get_tensor().set(np_array, place)
    m1o, m2o, po = (np.zeros((height, row_numel)) for _ in range(3))
    for idx, row_id in enumerate(rows):
        m1o[row_id] = beta1 * dense_inputs["Moment1"][row_id] + (1 - beta1) * np_array[idx]
        m2o[row_id] = beta2 * dense_inputs["Moment2"][row_id] + (1 - beta2) * np.square(np_array[idx])
        lr_t = dense_inputs["LearningRate"] * np.sqrt(1 - dense_inputs["Beta2Pow"]) / (1 - dense_inputs["Beta1Pow"])
        po[row_id] = dense_inputs["Param"][row_id] - lr_t * (m1o[row_id] / (np.sqrt(m2o[row_id]) + epsilon))
    op_args = {}
    for key, arr in dense_inputs.items():
        scope.var(key).get_tensor().set(arr, place)
        op_args[key] = key
    op_args.update(Grad="Grad", ParamOut="Param", Moment1Out="Moment1", Moment2Out="Moment2",
                   beta1=beta1, beta2=beta2, epsilon=epsilon)
    core.Operator("adam", **op_args).run(scope, place)
    for key, want in (("Param", po), ("Moment1", m1o), ("Moment2", m2o)):
        actual = np.array(scope.var(key).get_tensor())
        for row_id in rows:
            assert (np.abs((actual[row_id] - want[row_id]) / actual[row_id]) < 0.00001).all()
        untouched = [r for r in range(height) if r not in rows]
        assert (actual[untouched] == 5.0).all()


def test_selected_rows_sum_op_with_empty_inputs(core):
    # TestSelectedRowsSumOp (test_sum_op.py:47-105): any subset of the 3 inputs empty
    place = core.CUDAPlace(0)
    for has in [(1, 1, 1), (0, 1, 1), (0, 0, 1), (0, 0, 0), (1, 0, 1)]:
        scope = core.Scope()
        for name, h in zip(("W1", "W2", "W3"), has):
            rows = [0, 1, 2, 3, 4, 5, 6] if h else []
            w = scope.var(name).get_selected_rows()
            w.set_height(len(rows))
            w.set_rows(rows)
            arr = np.ones((len(rows), 12)).astype("float32")
            for i in range(len(rows)):
                arr[i] *= i
            w.get_tensor().set(arr, place)
        out = scope.var('Out').get_selected_rows()
        out.set_height(7 if has[0] else 0)
        if len(set(has)) == 1 or True:
            # heights must agree among non-empty inputs (sum_op.h:139); empty ones carry height 0
            out.set_height(7)
        for name, h in zip(("W1", "W2", "W3"), has):
            if not h:
                scope.var(name).get_selected_rows().set_height(7)
        core.Operator("sum", X=["W1", "W2", "W3"], Out='Out').run(scope, place)
        assert 7 * sum(has) == len(out.rows())
        if sum(has):
            assert out.rows() == [0, 1, 2, 3, 4, 5, 6] * sum(has)


def test_sum_into_dense_accumulates_selected_rows(core):
    # sum_op.h:40-70: LoDTensor output with a dense and two SelectedRows inputs (duplicates accumulate)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    dense = np.full((10, 10), 3.0, np.float32)
    scope.var("D").get_tensor().set(dense, place)
    for name, rows, fill in (("S1", [0, 4, 7], 1.0), ("S2", [0, 5, 7, 9], 2.0)):
        s = scope.var(name).get_selected_rows()
        s.set_height(10)
        s.set_rows(rows)
        s.get_tensor().set(np.full((len(rows), 10), fill, np.float32), place)
    scope.var("Out").get_tensor().set(np.zeros((10, 10), np.float32), place)
    core.Operator("sum", X=["D", "S1", "S2"], Out="Out").run(scope, place)
    out = np.array(scope.var("Out").get_tensor())
    # the expectations of selected_rows_functor_test.cc:100-120
    assert out[0, 0] == 6.0 and out[1, 1] == 3.0 and out[4, 6] == 4.0 and out[5, 7] == 5.0
    assert out[6, 1] == 3.0 and out[7, 3] == 6.0 and out[9, 6] == 5.0


def test_sparse_adagrad_op(core):
    # TestSparseAdagradOp.check_with_place (test_adagrad_op.py:87-188) — rows [0, 4, 7, 4], verbatim values
    import math
    place = core.CUDAPlace(0)
    scope = core.Scope()
    height, rows, row_numel = 10, [0, 4, 7, 4], 12
    grad_selected_rows = scope.var('Grad').get_selected_rows()
    grad_selected_rows.set_height(height)
    grad_selected_rows.set_rows(rows)
    np_array = np.ones((len(rows), row_numel)).astype("float32")
    np_array[0, 0] = 2.0
    np_array[2, 8] = 4.0
    grad_selected_rows.get_tensor().set(np_array, place)
    scope.var('Param').get_tensor().set(np.full((height, row_numel), 5.0).astype("float32"), place)
    scope.var('LearningRate').get_tensor().set(np.full((1), 2.0).astype("float32"), place)
    scope.var('Moment').get_tensor().set(np.full((height, row_numel), 2.0).astype("float32"), place)
    adagrad_op = core.Operator("adagrad", Param='Param', Grad='Grad', ParamOut='Param', Moment='Moment',
                               MomentOut='Moment', LearningRate='LearningRate', epsilon=2.0)
    adagrad_op.run(scope, place)
    moment_result_array = np.array(scope.var('Moment').get_tensor())
    assert moment_result_array[rows[0], 0] == pytest.approx(6.0)
    assert moment_result_array[rows[0], 2] == pytest.approx(3.0)
    assert moment_result_array[1, 0] == pytest.approx(2.0)
    assert moment_result_array[rows[1], 10] == pytest.approx(6.0)      # 2.0 + (1.0 + 1.0)^2
    assert moment_result_array[rows[3], 4] == pytest.approx(6.0)
    assert moment_result_array[5, 8] == pytest.approx(2.0)
    assert moment_result_array[rows[2], 1] == pytest.approx(3.0)
    assert moment_result_array[rows[2], 8] == pytest.approx(18.0)
    result_array = np.array(scope.var('Param').get_tensor())

    def get_out(param, lr, grad, m, epsilon):
        return param - lr * grad / (math.sqrt(m) + epsilon)
    assert result_array[rows[0], 0] == pytest.approx(get_out(5.0, 2.0, 2.0, 6.0, 2.0), abs=1e-5)
    assert result_array[rows[0], 2] == pytest.approx(get_out(5.0, 2.0, 1.0, 3.0, 2.0), abs=1e-5)
    assert result_array[1, 0] == pytest.approx(get_out(5.0, 2.0, 0.0, 2.0, 2.0), abs=1e-5)
    assert result_array[rows[1], 10] == pytest.approx(get_out(5.0, 2.0, 2.0, 6.0, 2.0), abs=1e-5)   # grad_merge = 1 + 1
    assert result_array[5, 8] == pytest.approx(get_out(5.0, 2.0, 0.0, 2.0, 2.0), abs=1e-5)
    assert result_array[rows[2], 1] == pytest.approx(get_out(5.0, 2.0, 1.0, 3.0, 2.0), abs=1e-5)
    assert result_array[rows[2], 8] == pytest.approx(get_out(5.0, 2.0, 4.0, 18.0, 2.0), abs=1e-5)


def test_sharding_kernels_are_registered(core):
    # CPU-only (`PADDLE_THROW` on GPU) or one-memcpy-per-row in the reference; CUDA kernels here
    for op in ("split_ids", "merge_ids", "split_selected_rows", "extract_rows"):
        assert core.registered_kernels(op) == 2


def test_split_ids_op_lod_tensor(core):
    # TestSplitIdsOp (test_split_ids_op.py:25-37)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    ids = np.array([[0], [2], [2], [3], [5], [5], [6]]).astype("int64")
    scope.var("Ids").get_tensor().set(ids, place)
    outs = ["out0", "out1", "out2"]
    for name in outs:
        scope.var(name).get_tensor()
    core.Operator("split_ids", Ids="Ids", Out=outs).run(scope, place)
    want = [[[0], [3], [6]], [], [[2], [2], [5], [5]]]
    for name, w in zip(outs, want):
        got = np.array(scope.var(name).get_tensor())
        assert got.shape == (len(w), 1) and got.reshape(-1).tolist() == [x[0] for x in w]
    # a larger batch against the oracle
    big = np.random.default_rng(0).integers(0, 1 << 31, (50000, 1)).astype("int64")
    scope.var("Ids").get_tensor().set(big, place)
    outs = ["s%d" % i for i in range(7)]
    for name in outs:
        scope.var(name).get_tensor()
    core.Operator("split_ids", Ids="Ids", Out=outs).run(scope, place)
    for name, w in zip(outs, oracle.split_ids(big.reshape(-1), 7)):
        assert np.array_equal(np.array(scope.var(name).get_tensor()).reshape(-1), w)


def test_split_ids_op_selected_rows(core):
    # TestSpliteIds (test_split_ids_op.py:40-87), run 3 times like the reference test
    place = core.CUDAPlace(0)
    scope = core.Scope()
    rows, height, row_numel = [0, 5, 7, 4, 9], 20, 2
    x = scope.var("X").get_selected_rows()
    x.set_rows(rows)
    x.set_height(height)
    arr = np.array([[r + j for j in range(row_numel)] for r in rows], np.float32)
    x.get_tensor().set(arr, place)
    outs_name = ["out%d" % i for i in range(3)]
    outs = [scope.var(n).get_selected_rows() for n in outs_name]
    op = core.Operator("split_ids", Ids="X", Out=outs_name)
    for _ in range(3):
        op.run(scope, place)
        for out, expect in zip(outs, [[0, 9], [7, 4], [5]]):
            assert out.rows() == expect and out.height() == height
            got = np.array(out.get_tensor())
            for j, row in enumerate(expect):
                assert got[j, 0] == float(row) and got[j, 1] == float(row + 1)
    # duplicated row ids: every copy takes the LAST occurrence's value (id_to_index, split_ids_op.h:71-76)
    rows = [3, 1, 3, 4, 1, 3]
    x.set_rows(rows)
    arr = np.arange(12, dtype=np.float32).reshape(6, 2)
    x.get_tensor().set(arr, place)
    op.run(scope, place)
    for out, (wr, wv) in zip(outs, oracle.split_ids_selected_rows(rows, arr, 3)):
        assert out.rows() == wr.tolist()
        assert np.array_equal(np.array(out.get_tensor()).reshape(-1, 2), wv)


def test_merge_ids_op(core):
    # TestMergeIdsOp (test_merge_ids_op.py:22-34)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    ids = np.array([[0], [2], [2], [3], [5], [5], [6]]).astype("int64")
    x0 = np.array([[0.1, 0.2], [0.2, 0.3], [0.3, 0.4]]).astype("float32")
    x1 = np.zeros((0, 2), "float32")
    x2 = np.array([[0.4, 0.5], [0.4, 0.5], [0.5, 0.6], [0.5, 0.6]]).astype("float32")
    want = np.array([[0.1, 0.2], [0.4, 0.5], [0.4, 0.5], [0.2, 0.3], [0.5, 0.6], [0.5, 0.6], [0.3, 0.4]], "float32")
    scope.var("Ids").get_tensor().set(ids, place)
    for name, x in (("x0", x0), ("x1", x1), ("x2", x2)):
        scope.var(name).get_tensor().set(x, place)
    core.Operator("merge_ids", Ids="Ids", X=["x0", "x1", "x2"], Out="Out").run(scope, place)
    assert np.array_equal(np.array(scope.var("Out").get_tensor()), want)
    # a shard with the wrong number of rows: "after merge, all data in x_tensor should be used"
    scope.var("x2").get_tensor().set(x2[:3], place)
    scope.var("x0").get_tensor().set(np.concatenate([x0, x0[:1]]), place)
    with pytest.raises(core.FluidLiteError):
        core.Operator("merge_ids", Ids="Ids", X=["x0", "x1", "x2"], Out="Out").run(scope, place)


def test_split_selected_rows_op(core):
    # TestSpliteSelectedRows.check_with_place / check_grad_with_place (test_split_selected_rows_op.py:36-126)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    rows, height, row_numel = [0, 5, 7, 4, 20], 20, 2
    x = scope.var("X").get_selected_rows()
    x.set_rows(rows)
    x.set_height(height)
    arr = np.ones((len(rows), row_numel)).astype("float32")
    arr[0, 0], arr[2, 1], arr[4, 1] = 2.0, 4.0, 8.0
    x.get_tensor().set(arr, place)
    height_sections = [5, 5, 5, 5, 3]
    outs_name = ["out%d" % i for i in range(len(height_sections))]
    outs = [scope.var(n).get_selected_rows() for n in outs_name]
    core.Operator("split_selected_rows", X="X", Out=outs_name, height_sections=height_sections).run(scope, place)
    assert outs[0].rows() == [0, 4] and outs[1].rows() == [0, 2] and outs[4].rows() == [0]
    assert outs[2].rows() == [] and outs[3].rows() == []
    assert outs[0].height() == height_sections[0] and outs[4].height() == height_sections[4]
    assert np.array(outs[0].get_tensor())[0, 0] == 2.0
    assert np.array(outs[1].get_tensor())[1, 1] == 4.0
    assert np.array(outs[4].get_tensor())[0, 1] == 8.0
    # its gradient is `sum` over the outputs' gradients (split_selected_rows_op.cc:64-77)
    g0 = scope.var("out0@GRAD").get_selected_rows()
    g0.set_rows([0, 5])
    g0.set_height(10)
    a0 = np.ones((2, row_numel)).astype("float32")
    a0[0, 0] = 2.0
    g0.get_tensor().set(a0, place)
    g1 = scope.var("out1@GRAD").get_selected_rows()
    g1.set_rows([2, 0])
    g1.set_height(10)
    a1 = np.ones((2, row_numel)).astype("float32")
    a1[0, 1] = 4.0
    g1.get_tensor().set(a1, place)
    xg = scope.var("X@GRAD").get_selected_rows()
    core.Operator("sum", X=["out0@GRAD", "out1@GRAD"], Out="X@GRAD").run(scope, place)
    assert xg.rows() == [0, 5, 2, 0] and xg.height() == 10
    got = np.array(xg.get_tensor())
    assert got[0, 0] == 2.0 and got[2, 1] == 4.0


def test_extract_rows_op(core):
    # TestExtractRows (test_extract_rows_op.py:26-53)
    place = core.CUDAPlace(0)
    scope = core.Scope()
    rows, feature_len = [0, 4, 4, 7], 12
    x = scope.var("X").get_selected_rows()
    x.set_height(len(rows))
    x.set_rows(rows)
    x.get_tensor().set(np.ones((len(rows), feature_len)).astype("float32"), place)
    scope.var("Out").get_tensor()
    core.Operator("extract_rows", X="X", Out="Out").run(scope, place)
    result = np.array(scope.var("Out").get_tensor())
    assert result.shape == (4, 1) and [e[0] for e in result] == rows


# ------------------------------------------------------------------ SparseAllToAllOpHandle
def _run_sharded_handles(core, devices, V, D, ids_per_dev, grads_per_dev, table, fused, steps=1, lr=0.5):
    """ShardedLookupOpHandle -> lookup_table_grad per device -> SparseAllToAllOpHandle, `steps` times."""
    G = len(devices)
    hs = (V + G - 1) // G
    pe = core.ShardedParallelSession(devices)
    for d in range(G):
        shard = np.zeros((hs, D), np.float32)
        rows = np.arange(d, V, G)
        shard[:len(rows)] = table[rows]
        sc, place = pe.scopes[d], core.CUDAPlace(devices[d])
        sc.var("emb_w").get_tensor().set(shard, place)
        sc.var("lr").get_tensor().set(np.array([lr], np.float32), place)
    n_max = max(len(i) for i in ids_per_dev)
    pe.shard_table("emb_w", V, D, n_max, 2 * n_max * G)
    core.set_fused(fused)
    outs = []
    try:
        for _ in range(steps):
            for d in range(G):
                pe.scopes[d].var("ids").get_tensor().set(ids_per_dev[d].reshape(-1, 1), core.CUDAPlace(devices[d]))
            pe.sharded_lookup("ids", "emb")
            assert pe.sync() == 0
            outs = [np.array(pe.scopes[d].var("emb").get_tensor()) for d in range(G)]
            for d in range(G):        # the per-device gradient op of the replica, as ParallelExecutor would run it
                sc, place = pe.scopes[d], core.CUDAPlace(devices[d])
                sc.var("emb@GRAD").get_tensor().set(grads_per_dev[d], place)
                sc.var("emb_w@GRAD").get_selected_rows()
                core.Operator("lookup_table_grad", W="emb_w", Ids="ids", is_sparse=True,
                              **{"Out@GRAD": "emb@GRAD", "W@GRAD": "emb_w@GRAD"}).run(sc, place)
            pe.sparse_all_to_all("emb_w@GRAD", "lr")
            assert pe.sync() == 0
    finally:
        core.set_fused(False)
    shards = [np.array(pe.scopes[d].var("emb_w").get_tensor()) for d in range(G)]
    full = np.stack(shards, axis=1).reshape(G * hs, D)[:V]
    return outs, full


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("fused", [False, True], ids=["materialised", "fused"])
def test_sparse_all_to_all_op_handle_reduce_test_case(core, world, fused):
    """reduce_op_handle_test.cc:153-209 (TestReduceSelectedRows) re-expressed for the sharded result: every
    device holds the SAME 20-row SelectedRows with duplicates, value[k] = k; where the reference checks the
    gathered rows == the concatenation in device order, here the table rows must have received exactly that
    concatenation: row r changes by lr * (sum over devices and occurrences of the value rows of r)."""
    kDims = (20, 20)
    V, D = kDims[0] * 2, kDims[1]
    rows = np.array([0, 1, 2, 3, 3, 0, 14, 7, 3, 1, 2, 4, 6, 3, 1, 1, 1, 1, 3, 7], np.int64)
    send = np.arange(kDims[0] * kDims[1], dtype=np.float32).reshape(kDims)
    table = np.random.default_rng(1).random((V, D)).astype(np.float32)
    devices = [0] * world if torch.cuda.device_count() < world else list(range(world))
    outs, full = _run_sharded_handles(core, devices, V, D, [rows] * world, [send] * world, table, fused)
    for o in outs:
        assert np.array_equal(o, table[rows])                           # sharded lookup == lookup_table on the full table
    want = table.copy()
    cat_rows, cat_val = np.tile(rows, world), np.tile(send, (world, 1))  # gather-concat in device order
    oracle.sgd_sparse(want, cat_rows, cat_val, np.array([0.5], np.float32))
    truth = table.astype(np.float64)
    r_, merged = oracle.merge_add(cat_rows, cat_val, truth=True)
    truth[r_] -= 0.5 * merged
    assert np.allclose(full, truth, rtol=1e-6, atol=1e-4)                # values up to 399 * lr * occurrences
    assert np.allclose(full, want, rtol=1e-5, atol=1e-3)
    untouched = np.setdiff1d(np.arange(V), rows)
    assert np.array_equal(full[untouched], table[untouched])


def test_sparse_all_to_all_op_handle_random_steps(core):
    """Three consecutive steps of different per-device batches on 3 ranks (sharing GPU 0 when the box has fewer)."""
    V, D, G = 5003, 32, 3
    rng = np.random.default_rng(2)
    table = (rng.random((V, D)) * 0.1 - 0.05).astype(np.float32)
    ids = [(rng.zipf(1.2, 4000 + 100 * d) % V).astype(np.int64) for d in range(G)]
    grads = [(rng.standard_normal((len(i), D)) * 0.01).astype(np.float32) for i in ids]
    devices = [0] * G if torch.cuda.device_count() < G else list(range(G))
    outs, full = _run_sharded_handles(core, devices, V, D, ids, grads, table, True, steps=3, lr=0.05)
    truth = table.astype(np.float64)
    bound = np.zeros_like(truth)
    all_i, all_g = np.concatenate(ids), np.concatenate(grads)
    r_, merged = oracle.merge_add(all_i, all_g, truth=True)
    _, l1 = oracle.merge_add(all_i, np.abs(all_g), truth=True)
    for _ in range(3):
        last = truth.copy()
        truth[r_] -= 0.05 * merged
        bound[r_] += 1e-6 * 0.05 * l1
    for d in range(G):      # the last lookup saw two updates
        assert np.abs(outs[d] - last[ids[d]]).max() < 1e-6
    assert (np.abs(full - truth) <= bound + 1e-6 * np.abs(truth) + 1e-30).all()
