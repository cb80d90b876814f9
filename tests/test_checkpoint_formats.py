"""Reference on-disk formats (golden G10): byte layout of LoDTensor / SelectedRows records
(framework/lod_tensor.cc:251-276, tensor_util.cc:244-290, selected_rows.cc:66-116) and round trips, incl. a
2-way "sharded" save/load done in one process."""
import io
import struct

import numpy as np
import pytest


@pytest.fixture(scope="module")
def ck(pkg):
    import importlib
    return importlib.import_module("paddle1_b200.checkpoint")


def test_tensor_desc_protobuf_bytes(ck):
    # data_type FP32 = 5 (framework.proto:105), dims [10, 100]: 08 05 | 10 0a | 10 64
    assert ck.encode_tensor_desc(np.float32, (10, 100)) == bytes([0x08, 0x05, 0x10, 0x0A, 0x10, 0x64])
    # negative dim (-1 = unknown) is a 10-byte varint, like protobuf's int64
    enc = ck.encode_tensor_desc(np.int64, (-1, 640))
    assert enc[:2] == bytes([0x08, 0x03]) and enc[2] == 0x10 and enc[3:13] == b"\xff" * 9 + b"\x01"
    assert ck.decode_tensor_desc(enc) == (np.dtype(np.int64), [-1, 640])
    # packed dims are accepted when reading
    assert ck.decode_tensor_desc(bytes([0x08, 0x06, 0x12, 0x02, 0x0A, 0x64])) == (np.dtype(np.float64), [10, 100])


def test_lod_tensor_layout_and_round_trip(ck):
    a = np.arange(12, dtype=np.float32).reshape(3, 4)
    raw = ck.dumps_lod_tensor(a, lod=[[0, 1, 3]])
    assert struct.unpack_from("<I", raw, 0) == (0,)                 # LoDTensor version
    assert struct.unpack_from("<Q", raw, 4) == (1,)                 # one LoD level
    assert struct.unpack_from("<Q", raw, 12) == (24,)               # 3 size_t offsets
    assert struct.unpack_from("<QQQ", raw, 20) == (0, 1, 3)
    assert struct.unpack_from("<I", raw, 44) == (0,)                # Tensor version
    (desc_size,) = struct.unpack_from("<i", raw, 48)
    assert raw[52:52 + desc_size] == ck.encode_tensor_desc(np.float32, (3, 4))
    assert raw[52 + desc_size:] == a.tobytes()
    b, lod = ck.read_lod_tensor(io.BytesIO(raw))
    assert np.array_equal(a, b) and lod == [[0, 1, 3]]


def test_selected_rows_round_trip_g10(ck, goldens):
    # framework/selected_rows_test.cc:20-58: rows {0,4,7}, height 10, value [3, 100]
    g = goldens["G10_selected_rows_layout"]
    value = np.random.default_rng(0).standard_normal((len(g["rows"]), g["row_numel"])).astype(np.float32)
    buf = io.BytesIO()
    ck.write_selected_rows(buf, g["rows"], g["height"], value)
    raw = buf.getvalue()
    assert struct.unpack_from("<I", raw, 0) == (0,)
    assert struct.unpack_from("<Q", raw, 4) == (3,)
    assert struct.unpack_from("<qqq", raw, 12) == (0, 4, 7)
    assert struct.unpack_from("<q", raw, 36) == (10,)
    rows, height, v = ck.read_selected_rows(io.BytesIO(raw))
    assert rows.tolist() == g["rows"] and height == g["height"] and np.array_equal(v, value)
    assert [height, v.shape[1]] == g["complete_dims"]                # GetCompleteDims


def test_sharded_save_load_is_global_row_order(ck, tmp_path):
    import torch
    v, d, world = 1003, 8, 1
    full = np.random.default_rng(1).standard_normal((v, d)).astype(np.float32)
    path = str(tmp_path / "table")
    ck.save_sharded(path, torch.from_numpy(full), v, 0, 1)
    with open(path, "rb") as f:
        back, lod = ck.read_lod_tensor(f)
    assert lod == [] and np.array_equal(back, full)
    # the gather-through-rank-0 writer produces the same bytes as the every-rank-writes-its-rows one
    ck.save_sharded(path + ".g", torch.from_numpy(full), v, 0, 1, parallel=False)
    assert open(path, "rb").read() == open(path + ".g", "rb").read()
    # a reader with 3 ranks takes rows r, r+3, ... (owner = id % G, operators/split_ids_op.h:47)
    for rank in range(3):
        shard, height = ck.load_sharded(path, rank, 3, "cpu")
        assert height == v
        mine = full[rank::3]
        assert np.array_equal(shard[: mine.shape[0]].numpy(), mine)
        assert (shard[mine.shape[0]:] == 0).all()
