"""Checkpoints of the embedding table and optimizer moments in the reference's on-disk formats,
so a table trained here loads with `fluid.io.load_persistables` and vice versa.

  LoDTensor   (framework/lod_tensor.cc:251-276 + tensor_util.cc:244-290):
      u32 version=0 | u64 lod_levels | per level: u64 bytes, size_t offsets[] |
      u32 version=0 | i32 desc_size | TensorDesc protobuf {1: data_type enum, 2: repeated int64 dims} | raw data
  SelectedRows (framework/selected_rows.cc:66-116):
      u32 version=0 | u64 nrows | i64 rows[] | i64 height | Tensor record (as above, from the 2nd u32 version)

The protobuf message is two scalar fields, encoded by hand here (the image has no protoc).  A row-sharded table
(owner = id % G) is written as ONE [V, D] LoDTensor in global row order, streamed in slabs through rank 0.
"""
import io
import struct

import numpy as np

# framework.proto:98-111
_PROTO_TYPE = {np.dtype(np.bool_): 0, np.dtype(np.int16): 1, np.dtype(np.int32): 2, np.dtype(np.int64): 3,
               np.dtype(np.float16): 4, np.dtype(np.float32): 5, np.dtype(np.float64): 6,
               np.dtype(np.uint8): 20, np.dtype(np.int8): 21}
_NP_TYPE = {v: k for k, v in _PROTO_TYPE.items()}


def _varint(x):
    x &= (1 << 64) - 1            # int64 as two's complement (negative dims take 10 bytes)
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _read_varint(buf, pos):
    x, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        x |= (b & 0x7F) << shift
        shift += 7
        if not b & 0x80:
            break
    if x >= 1 << 63:
        x -= 1 << 64
    return x, pos


def encode_tensor_desc(dtype, dims):
    """VarType.TensorDesc: `required Type data_type = 1; repeated int64 dims = 2;` (proto2: dims unpacked)."""
    out = b"\x08" + _varint(_PROTO_TYPE[np.dtype(dtype)])
    for d in dims:
        out += b"\x10" + _varint(int(d))
    return out


def decode_tensor_desc(buf):
    pos, dtype, dims = 0, None, []
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if field == 1 and wire == 0:
            v, pos = _read_varint(buf, pos)
            dtype = _NP_TYPE[v]
        elif field == 2 and wire == 0:
            v, pos = _read_varint(buf, pos)
            dims.append(v)
        elif field == 2 and wire == 2:       # packed encoding, accepted on read
            n, pos = _read_varint(buf, pos)
            end = pos + n
            while pos < end:
                v, pos = _read_varint(buf, pos)
                dims.append(v)
        else:
            raise ValueError("unexpected field %d (wire type %d) in TensorDesc" % (field, wire))
    return dtype, dims


def write_tensor_header(f, dtype, dims):
    desc = encode_tensor_desc(dtype, dims)
    f.write(struct.pack("<I", 0))
    f.write(struct.pack("<i", len(desc)))
    f.write(desc)


def write_tensor(f, array):
    """TensorToStream"""
    array = np.ascontiguousarray(array)
    write_tensor_header(f, array.dtype, array.shape)
    f.write(array.tobytes())


def read_tensor(f):
    """TensorFromStream -> ndarray"""
    (version,) = struct.unpack("<I", f.read(4))
    if version != 0:
        raise ValueError("Only version 0 is supported")
    (size,) = struct.unpack("<i", f.read(4))
    dtype, dims = decode_tensor_desc(f.read(size))
    n = int(np.prod(dims)) if dims else 1
    data = f.read(n * dtype.itemsize)
    return np.frombuffer(data, dtype=dtype).reshape(dims).copy()


def write_lod_header(f, lod=()):
    f.write(struct.pack("<I", 0))
    f.write(struct.pack("<Q", len(lod)))
    for level in lod:
        level = np.asarray(level, dtype=np.uint64)
        f.write(struct.pack("<Q", level.size * 8))
        f.write(level.tobytes())


def write_lod_tensor(f, array, lod=()):
    """SerializeToStream(LoDTensor)"""
    write_lod_header(f, lod)
    write_tensor(f, array)


def read_lod_tensor(f):
    """DeserializeFromStream(LoDTensor) -> (ndarray, lod)"""
    (version,) = struct.unpack("<I", f.read(4))
    if version != 0:
        raise ValueError("Only version 0 is supported")
    (levels,) = struct.unpack("<Q", f.read(8))
    lod = []
    for _ in range(levels):
        (nbytes,) = struct.unpack("<Q", f.read(8))
        lod.append(np.frombuffer(f.read(nbytes), dtype=np.uint64).astype(np.int64).tolist())
    return read_tensor(f), lod


def write_selected_rows(f, rows, height, value):
    """SerializeToStream(SelectedRows)"""
    rows = np.ascontiguousarray(rows, dtype=np.int64)
    f.write(struct.pack("<I", 0))
    f.write(struct.pack("<Q", rows.size))
    f.write(rows.tobytes())
    f.write(struct.pack("<q", int(height)))
    write_tensor(f, value)


def read_selected_rows(f):
    """-> (rows, height, value)"""
    (version,) = struct.unpack("<I", f.read(4))
    if version != 0:
        raise ValueError("Only version 0 is supported")
    (n,) = struct.unpack("<Q", f.read(8))
    rows = np.frombuffer(f.read(8 * n), dtype=np.int64).copy()
    (height,) = struct.unpack("<q", f.read(8))
    return rows, height, read_tensor(f)


def dumps_lod_tensor(array, lod=()):
    b = io.BytesIO()
    write_lod_tensor(b, array, lod)
    return b.getvalue()


# ------------------------------------------------------------------ row-sharded tables
def save_sharded(path, shard, height, rank, world, group=None, slab_rows=1 << 20, parallel=True):
    """Write the full [height, D] table (global row r lives at shard[r // world] of rank r % world) as one
    reference-format LoDTensor file that fluid.io.load_persistables can read.  Every rank calls this.
    parallel=True (ranks share a file system): rank 0 writes the header and sizes the file, then EVERY rank
    stores its own rows (rank, rank + world, ...) through a shared memory map, slab by slab — no rank ever
    holds more than one slab of someone else's data and nothing funnels through rank 0.
    parallel=False: slabs are all-gathered and rank 0 writes them (works without a shared file system)."""
    import torch
    import torch.distributed as dist
    d = shard.shape[1]
    np_dtype = np.dtype(str(shard.dtype).replace("torch.", ""))
    hs = shard.shape[0]
    multi = world > 1 and dist.is_available() and dist.is_initialized()
    if parallel:
        offset = 0
        if rank == 0:
            with open(path, "wb") as f:
                write_lod_header(f)
                write_tensor_header(f, np_dtype, (height, d))
                offset = f.tell()
                f.truncate(offset + height * d * np_dtype.itemsize)
        if multi:
            box = [offset]
            dist.broadcast_object_list(box, src=0, group=group)
            offset = box[0]
        full = np.memmap(path, dtype=np_dtype, mode="r+", offset=offset, shape=(height, d))
        mine = full[rank::world]                      # this rank's rows, strided view of the file
        for lo in range(0, mine.shape[0], slab_rows):
            hi = min(mine.shape[0], lo + slab_rows)
            mine[lo:hi] = shard[lo:hi].cpu().numpy()
        full.flush()
        del mine, full
        if multi:
            dist.barrier(group=group)
        return
    f = None
    if rank == 0:
        f = open(path, "wb")
        write_lod_header(f)
        write_tensor_header(f, np_dtype, (height, d))
    written = 0
    for lo in range(0, hs, slab_rows):
        part = shard[lo:lo + slab_rows].contiguous()
        if world > 1:
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part, group=group)
            block = torch.stack(parts, dim=1).reshape(-1, d)          # interleave: global row order
        else:
            block = part
        if rank == 0:
            take = min(block.shape[0], height - written)
            f.write(block[:take].cpu().numpy().tobytes())
            written += take
    if rank == 0:
        f.close()


def load_sharded(path, rank, world, device, dtype=None):
    """Read this rank's rows (rank, rank + world, ...) of a reference-format [V, D] LoDTensor file."""
    import torch
    with open(path, "rb") as f:
        (version,) = struct.unpack("<I", f.read(4))
        (levels,) = struct.unpack("<Q", f.read(8))
        if version != 0:
            raise ValueError("Only version 0 is supported")
        for _ in range(levels):
            (nbytes,) = struct.unpack("<Q", f.read(8))
            f.seek(nbytes, 1)
        struct.unpack("<I", f.read(4))
        (size,) = struct.unpack("<i", f.read(4))
        np_dtype, dims = decode_tensor_desc(f.read(size))
        offset = f.tell()
    full = np.memmap(path, dtype=np_dtype, mode="r", offset=offset, shape=tuple(dims))
    mine = np.ascontiguousarray(full[rank::world])
    hs = (dims[0] + world - 1) // world
    out = torch.zeros((hs, dims[1]), dtype=dtype or getattr(torch, str(np_dtype)), device=device)
    out[: mine.shape[0]] = torch.from_numpy(mine).to(device)
    return out, dims[0]
