#!/usr/bin/env python
"""Writes tests/golden/goldens.json: the known-answer vectors the reference's OWN tests hold
for the sparse-embedding path, transcribed as data.

The reference (`paddle`) cannot be imported here (no core.so), so these are the hand-written
constants of its tests, each with the file:line they come from under /root/reference.
Inputs are given in full; `expect` lists (index -> value) pairs exactly as asserted there.
Run:  python tests/golden/make_goldens.py
"""
import json
import os

G = {}

# G1/G2  paddle/fluid/operators/math/selected_rows_functor_test.cc:20-121 (cpu_add) and
# :123-221 (cpu_add_to): rows1={0,4,7} filled 1.0, rows2={0,5,7,9} filled 2.0, height 10,
# row_numel 10; concat; then + tensor filled 3.0.
G["G1_G2_selected_rows_add"] = {
    "src": "operators/math/selected_rows_functor_test.cc:20-221",
    "height": 10, "row_numel": 10,
    "rows1": [0, 4, 7], "fill1": 1.0,
    "rows2": [0, 5, 7, 9], "fill2": 2.0,
    "tensor_fill": 3.0,
    "expect_rows": [0, 4, 7, 0, 5, 7, 9],
    # (row, col) -> value of the concatenated value block
    "expect_concat": [[0, 0, 1.0], [0, 8, 1.0], [1, 1, 1.0], [2, 6, 1.0],
                      [3, 3, 2.0], [3, 8, 2.0], [4, 4, 2.0], [5, 7, 2.0], [6, 9, 2.0]],
    # (row, col) -> value of the dense result
    "expect_dense": [[0, 0, 6.0], [1, 1, 3.0], [4, 6, 4.0], [5, 7, 5.0], [6, 1, 3.0],
                     [7, 3, 6.0], [9, 6, 5.0]],
}

# G3  python/paddle/fluid/tests/unittests/test_sgd_op.py:38-99
G["G3_sparse_sgd"] = {
    "src": "python/paddle/fluid/tests/unittests/test_sgd_op.py:38-99",
    "height": 10, "rows": [0, 4, 7], "row_numel": 12,
    "grad_fill": 1.0, "grad_set": [[0, 0, 2.0], [2, 8, 4.0]],
    "param_fill": 5.0, "lr": 2.0,
    "expect": [[0, 0, 1.0], [0, 2, 3.0], [1, 0, 5.0], [4, 10, 3.0], [5, 8, 5.0],
               [7, 1, 3.0], [7, 8, -3.0]],
}

# G4  test_adam_op.py:232-313 with the numpy model adam_step_sparse (:197-229); rel 1e-5 on
# touched rows.
G["G4_sparse_adam"] = {
    "src": "python/paddle/fluid/tests/unittests/test_adam_op.py:197-313",
    "height": 10, "rows": [0, 4, 7], "row_numel": 12,
    "grad_fill": 1.0, "grad_set": [[0, 0, 2.0], [2, 8, 4.0]],
    "param_fill": 5.0, "moment1_fill": 5.0, "moment2_fill": 5.0, "lr": 2.0,
    "beta1": 0.78, "beta2": 0.836, "epsilon": 1e-4, "beta_pow_exponent": 10,
    "rel_tol": 1e-5,
}

# G5  test_adagrad_op.py:87-188 — rows [0,4,7,4]: the only duplicate-row MergeAdd golden.
G["G5_sparse_adagrad"] = {
    "src": "python/paddle/fluid/tests/unittests/test_adagrad_op.py:87-188",
    "height": 10, "rows": [0, 4, 7, 4], "row_numel": 12,
    "grad_fill": 1.0, "grad_set": [[0, 0, 2.0], [2, 8, 4.0]],
    "param_fill": 5.0, "moment_fill": 2.0, "lr": 2.0, "epsilon": 2.0,
    "expect_moment": [[0, 0, 6.0], [0, 2, 3.0], [1, 0, 2.0], [4, 10, 6.0], [4, 4, 6.0],
                      [5, 8, 2.0], [7, 1, 3.0], [7, 8, 18.0]],
    # (row, col, grad, moment) -> param = 5 - 2*grad/(sqrt(moment)+2), to 5 places
    "expect_param": [[0, 0, 2.0, 6.0], [0, 2, 1.0, 3.0], [1, 0, 0.0, 2.0], [4, 10, 2.0, 6.0],
                     [5, 8, 0.0, 2.0], [7, 1, 1.0, 3.0], [7, 8, 4.0, 18.0]],
}

# G6  test_lookup_table_op.py:25-83 — Out == table[ids]; table 17x31; ids (4,1) and (2,4,5,1);
# padding rows zero (padding_idx = random id from ids, negative padding_idx wraps by +17).
G["G6_lookup_table"] = {
    "src": "python/paddle/fluid/tests/unittests/test_lookup_table_op.py:25-83",
    "table_shape": [17, 31], "ids_shapes": [[4, 1], [2, 4, 5, 1]],
}

# G7  test_sum_op.py:47-105 — 3 SelectedRows inputs, rows [0..6], row_numel 12, value row i = i;
# any subset empty (1 = has data below); len(out.rows) == 7 * (#non-empty).
G["G7_sum_selected_rows"] = {
    "src": "python/paddle/fluid/tests/unittests/test_sum_op.py:47-105",
    "height": 7, "row_numel": 12, "rows": [0, 1, 2, 3, 4, 5, 6],
    "cases": [[1, 1, 1], [0, 1, 1], [1, 0, 1], [1, 1, 0], [0, 0, 1], [1, 0, 0], [0, 1, 0], [0, 0, 0]],
}

# G8  framework/details/reduce_op_handle_test.cc:153-209 — every device holds the same 20 rows
# (with duplicates), value = arange(20*20) reshaped to kDims {20,20}, height = 40; the reduced
# SelectedRows == concatenation in device order (rows[k % 20], value[j % 400]).
G["G8_gather_selected_rows"] = {
    "src": "framework/details/reduce_op_handle_test.cc:29,153-209",
    "height": 40, "width": 20,
    "rows": [0, 1, 2, 3, 3, 0, 14, 7, 3, 1, 2, 4, 6, 3, 1, 1, 1, 1, 3, 7],
}

# G9  test_split_ids_op.py:26-78, test_merge_ids_op.py:22-34, test_split_selected_rows_op.py:36-83
G["G9_split_merge_ids"] = {
    "src": "python/paddle/fluid/tests/unittests/test_{split_ids,merge_ids,split_selected_rows}_op.py",
    "ids": [0, 2, 2, 3, 5, 5, 6], "shard_num": 3,
    "expect_shards": [[0, 3, 6], [], [2, 2, 5, 5]],
    "sr_rows": [0, 5, 7, 4, 9], "sr_height": 20, "sr_row_numel": 2,
    "expect_sr_rows": [[0, 9], [7, 4], [5]],
    "merge_x": [[[0.1, 0.2], [0.2, 0.3], [0.3, 0.4]], [],
                [[0.4, 0.5], [0.4, 0.5], [0.5, 0.6], [0.5, 0.6]]],
    "merge_expect": [[0.1, 0.2], [0.4, 0.5], [0.4, 0.5], [0.2, 0.3], [0.5, 0.6], [0.5, 0.6], [0.3, 0.4]],
    "ssr_rows": [0, 5, 7, 4, 20], "ssr_row_numel": 2,
    "ssr_value_set": [[0, 0, 2.0], [2, 1, 4.0], [4, 1, 8.0]],
    "ssr_height_sections": [5, 5, 5, 5, 3],
    "ssr_expect_rows": {"0": [0, 4], "1": [0, 2], "4": [0]},
    "ssr_expect_value": [[0, 0, 0, 2.0], [1, 1, 1, 4.0], [4, 0, 1, 8.0]],
}

# G10  framework/selected_rows_test.cc:44-58 + selected_rows.cc:66-116 on-disk layout:
# u32 version(0), u64 nrows, i64 rows[], i64 height, then the Tensor record.
G["G10_selected_rows_layout"] = {
    "src": "framework/selected_rows_test.cc:20-58; framework/selected_rows.cc:66-116",
    "rows": [0, 4, 7], "height": 10, "row_numel": 100,
    "complete_dims": [10, 100],
}

# G11  the ops of the sparse regulariser chain (regularizer.py:146-170, 226-253):
# test_extract_rows_op.py:28-53 (rows with a duplicate come back verbatim, shape [n,1]);
# test_scale_op.py:38-78 (SelectedRows in/out: value * scale exactly, rows and height kept);
# test_sign_op.py:23-30 (Out = np.sign(X)).
G["G11_regulariser_ops"] = {
    "src": "python/paddle/fluid/tests/unittests/test_{extract_rows,scale,sign}_op.py",
    "extract_rows": [0, 4, 4, 7], "extract_feature_len": 12,
    "scale_rows": [0, 4, 7], "scale_height": 10, "scale_row_numel": 12, "scale": 2.0,
    "dense_scale": -2.3, "dense_shape": [10, 10],
    "sign_range": [-10, 10], "sign_shape": [10, 10],
}

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "goldens.json")
    with open(out, "w") as f:
        json.dump(G, f, indent=1, sort_keys=True)
    print("wrote", out)
