"""world_size-2 gloo test (CPU) of the row-sharded step's host logic (paddle-1_b200/sharded.py):
split sizes, owner convention id % G (operators/split_ids_op.h:45-49), merge_ids ordering
(operators/merge_ids_op.h:63-87) and the reference's gather-concat order for the bit-exact mode
(framework/details/reduce_and_gather.h:56-93).  Compute runs on the CPU checker backend
(tests/sharded_cpu_ops.py); the result must equal the single-table oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
V, D, N = 1003, 12, 700


def _inputs(rank):
    rng = np.random.default_rng(100 + rank)
    w = np.arange(1, V + 1, dtype=np.float64) ** -1.1
    cdf = np.cumsum(w) / w.sum()
    ids = ((np.searchsorted(cdf, rng.random(N)) * 2654435761 + 12345) % V).astype(np.int64)
    dout = (rng.standard_normal((N, D)) * 0.01).astype(np.float32)
    return ids, dout


def _table():
    return (np.random.default_rng(7).random((V, D)) * 0.1 - 0.05).astype(np.float32)


def _worker(rank, world, port, deterministic, results):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import __graft_entry__ as ge
        from sharded_cpu_ops import CpuCheckerOps
        sharded = ge.load_package().sharded
        full = _table()
        tab = sharded.ShardedTable(V, D, CpuCheckerOps(), torch.device("cpu"),
                                   init=lambda rows: torch.from_numpy(full[rows.numpy()]))
        ids, dout = _inputs(rank)
        ws = tab.ops.alloc_workspace(N, D, 4, "cpu")
        out, ctx = tab.lookup(torch.from_numpy(ids), ws, deterministic=deterministic)
        fwd_ok = np.array_equal(out.numpy(), full[ids])
        tab.update(ctx, torch.from_numpy(dout), torch.tensor([0.05]), deterministic=deterministic)
        after = tab.gather_full_table().numpy()
        # shard-aware checkpoint in the reference's LoDTensor format: one [V, D] file in global row order
        import importlib
        import tempfile
        ck = importlib.import_module("paddle1_b200.checkpoint")
        path = os.path.join(tempfile.gettempdir(), "b200emb_ckpt_%d" % port)
        ck.save_sharded(path, tab.shard, V, rank, world)
        dist.barrier()
        if rank == 0:
            with open(path, "rb") as f:
                saved, _ = ck.read_lod_tensor(f)
            fwd_ok = fwd_ok and np.array_equal(saved, after)
        back, _ = ck.load_sharded(path, rank, world, "cpu")
        fwd_ok = fwd_ok and torch.equal(back, tab.shard)
        dist.barrier()
        if rank == 0:
            os.remove(path)
            results.put((fwd_ok, after))
        else:
            results.put((fwd_ok, None))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("deterministic", [True, False], ids=["bit-exact", "fast"])
def test_two_rank_sharded_step_equals_single_table(deterministic):
    sys.path.insert(0, ROOT)
    import oracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, deterministic, q))
             for port in [_free_port()] for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for ok, _ in got), "forward rows differ from table[ids]"
    after = next(a for _, a in got if a is not None)
    # reference order: every device's (rows, grads) concatenated in device order, then applied
    ids = np.concatenate([_inputs(r)[0] for r in range(world)])
    dout = np.concatenate([_inputs(r)[1] for r in range(world)])
    want = _table()
    lr = np.array([0.05], np.float32)
    if deterministic:
        oracle.sgd_sparse(want, ids, dout, lr)
        assert np.array_equal(after.view(np.uint32), want.view(np.uint32))
    else:
        rows, merged = oracle.merge_add(ids, dout, truth=True)
        truth = want.astype(np.float64)
        truth[rows] -= 0.05 * merged
        _, l1 = oracle.merge_add(ids, np.abs(dout), truth=True)
        bound = np.zeros_like(truth)
        bound[rows] = 1e-6 * 0.05 * l1
        assert (np.abs(after - truth) <= bound + 1e-6 * np.abs(want) + 1e-30).all()
        untouched = np.setdiff1d(np.arange(V), ids)
        assert np.array_equal(after[untouched], want[untouched])
