"""GPU parity of the §8f "next" rows built on the hot path's kernels: split_selected_rows, extract_rows and
the fused sparse weight-decay chain (extract_rows -> lookup_table -> [sign] -> scale -> elementwise_add).
Everything is compared bit for bit with the oracle (byte/index work and separately rounded fp32/fp64 ops)."""
import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg.capi.lib()
    return pkg.capi


def _bits(a):
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


def _same(got, want):
    """bit-identical, except that a NaN only has to be a NaN (the GPU returns the canonical quiet NaN)"""
    nan = np.isnan(want)
    return np.array_equal(np.isnan(got), nan) and np.array_equal(_bits(got)[~nan], _bits(want)[~nan])


def _check_split(capi, rows, value, sections):
    want = oracle.split_selected_rows(rows, value, sections)
    out_rows, out_value, bounds = capi.split_selected_rows(
        torch.from_numpy(np.asarray(rows, np.int64)).cuda(), torch.from_numpy(value).cuda(), sections)
    b = bounds.cpu().numpy()
    assert b[0] == 0 and b[-1] == len(rows)
    for s, (wr, wv) in enumerate(want):
        assert out_rows[b[s]:b[s + 1]].cpu().numpy().tolist() == wr.tolist()
        assert np.array_equal(_bits(out_value[b[s]:b[s + 1]].cpu().numpy()), _bits(wv))


def test_split_selected_rows_golden(capi, goldens):
    # test_split_selected_rows_op.py:36-83
    g = goldens["G9_split_merge_ids"]
    v = np.full((len(g["ssr_rows"]), g["ssr_row_numel"]), 1.0, np.float32)
    for r, c, x in g["ssr_value_set"]:
        v[r, c] = x
    _check_split(capi, g["ssr_rows"], v, g["ssr_height_sections"])
    out_rows, out_value, bounds = capi.split_selected_rows(
        torch.tensor(g["ssr_rows"], dtype=torch.int64).cuda(), torch.from_numpy(v).cuda(), g["ssr_height_sections"])
    b = bounds.cpu().numpy()
    for k, want in g["ssr_expect_rows"].items():
        assert out_rows[b[int(k)]:b[int(k) + 1]].cpu().tolist() == want
    for sec, r, c, val in g["ssr_expect_value"]:
        assert out_value[b[sec] + r, c].item() == val


@pytest.mark.parametrize("n,d,sections", [
    (1, 1, [7]), (1000, 12, [100, 0, 300, 50]), (70000, 64, [1 << 14] * 8), (250000, 8, [5, 999990, 5]),
    (4096, 100, [1] * 64),
])
def test_split_selected_rows_random(capi, n, d, sections):
    rng = np.random.default_rng(n + d)
    total = int(sum(sections))
    rows = rng.integers(0, total + 3, n)          # a few rows past the last section start: last section takes them
    value = rng.standard_normal((n, d)).astype(np.float32)
    _check_split(capi, rows, value, sections)


def test_split_selected_rows_edges(capi):
    z = torch.empty((0, 4), device="cuda")
    r, v, b = capi.split_selected_rows(torch.empty(0, dtype=torch.int64, device="cuda"), z, [3, 3])
    assert r.numel() == 0 and b.cpu().tolist() == [0, 0, 0]
    rows = torch.arange(10, device="cuda")
    val = torch.ones((10, 4), device="cuda")
    with pytest.raises(capi.B200EmbError):
        capi.split_selected_rows(rows, val, [1] * 65)            # more than 64 sections
    with pytest.raises(capi.B200EmbError):
        capi.split_selected_rows(rows, val, [4, -1])
    # float64 values
    v64 = np.random.default_rng(0).standard_normal((10, 3))
    _, ov, bb = capi.split_selected_rows(rows, torch.from_numpy(v64).cuda(), [5, 5])
    assert np.array_equal(ov.cpu().numpy(), v64) and bb.cpu().tolist() == [0, 5, 10]


def test_extract_rows_golden(capi, goldens):
    g = goldens["G11_regulariser_ops"]
    rows = torch.tensor(g["extract_rows"], dtype=torch.int64).cuda()
    out = capi.extract_rows(rows)
    assert out.shape == (4, 1) and out.reshape(-1).cpu().tolist() == g["extract_rows"]
    assert capi.extract_rows(torch.empty(0, dtype=torch.int64, device="cuda")).numel() == 0


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("n,v,d", [(4, 8, 12), (1, 1, 1), (5000, 777, 7), (106496, 100000, 64), (30000, 5000, 100),
                                   (2000, 300, 520)])
def test_sparse_decay_matches_op_chain(capi, dtype, mode, n, v, d):
    rng = np.random.default_rng(n * 7 + d + mode)
    table = rng.standard_normal((v, d)).astype(dtype)
    table[0, :min(d, 3)] = [0.0, -0.0, np.nan][:min(d, 3)]
    rows = rng.integers(0, v, n)
    rows[0] = 0
    grad = rng.standard_normal((n, d)).astype(dtype)
    coeff = -0.0123 if mode else 0.0456
    t_d, r_d, g_d = torch.from_numpy(table).cuda(), torch.from_numpy(rows).cuda(), torch.from_numpy(grad).cuda()
    status = capi.new_status()
    decay = capi.sparse_decay(t_d, r_d, coeff, mode, status=status)
    assert _same(decay.cpu().numpy(), oracle.sparse_decay(table, rows, coeff, mode))
    want = oracle.sparse_decay(table, rows, coeff, mode, addend=grad)
    out = capi.sparse_decay(t_d, r_d, coeff, mode, addend=g_d, status=status)
    assert _same(out.cpu().numpy(), want)
    capi.sparse_decay(t_d, r_d, coeff, mode, addend=g_d, out=g_d, status=status)      # in place on the gradient
    assert _same(g_d.cpu().numpy(), want)
    assert int(status.item()) == 0


def test_sparse_decay_golden_and_bad_rows(capi, goldens):
    g = goldens["G11_regulariser_ops"]
    rng = np.random.default_rng(11)
    rows = g["scale_rows"]
    table = np.zeros((g["scale_height"], g["scale_row_numel"]), np.float32)
    in_array = rng.random((len(rows), g["scale_row_numel"])).astype(np.float32)
    table[rows] = in_array
    got = capi.sparse_decay(torch.from_numpy(table).cuda(), torch.tensor(rows).cuda(), g["scale"])
    assert (got.cpu().numpy() == in_array * np.float32(g["scale"])).all()       # test_scale_op.py:71
    # a row outside the table: flagged, no decay added (the reference's lookup_table would abort)
    status = capi.new_status()
    add = torch.ones((2, g["scale_row_numel"]), device="cuda")
    out = capi.sparse_decay(torch.from_numpy(table).cuda(), torch.tensor([4, 10]).cuda(), 2.0, addend=add,
                            status=status)
    assert int(status.item()) & 1
    assert torch.equal(out[1], add[1])
    with pytest.raises(capi.B200EmbError):
        capi.sparse_decay(torch.from_numpy(table).cuda(), torch.tensor([1]).cuda(), 2.0, mode=2)
