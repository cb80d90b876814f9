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
    for op in ("lookup_table", "lookup_table_grad", "sum", "sgd", "adam", "adagrad"):
        assert core.registered_kernels(op) == 2        # float + double, CUDAPlace


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
