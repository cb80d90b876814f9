"""GPU tests of the sharded-path kernels (csrc/shard.cu + key-mapped sort plan) through the C ABI,
and of ShardedTable on the CUDA backend: world 1 always, world 2 over NCCL when two GPUs exist."""
import os
import socket
import sys

import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


@pytest.fixture(scope="module")
def capi(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg.capi.lib()
    return pkg.capi


def _zipf_ids(rng, n, v):
    w = np.arange(1, v + 1, dtype=np.float64) ** -1.1
    cdf = np.cumsum(w) / w.sum()
    return ((np.searchsorted(cdf, rng.random(n)) * 2654435761 + 12345) % v).astype(np.int64)


@pytest.mark.parametrize("n,v,shards", [(5000, 777, 1), (5000, 1003, 4), (150000, 100000, 8), (300000, 4000000, 3)])
def test_plan_unique_rows_matches_split_ids_semantics(capi, n, v, shards):
    from sharded_cpu_ops import CpuCheckerOps, _Plan
    rng = np.random.default_rng(n + shards)
    ids = _zipf_ids(rng, n, v)
    ids[::97] = v + 5            # a few out-of-range ids: flagged and dropped
    d_ids = torch.from_numpy(ids).cuda()
    ws = capi.alloc_workspace(n, 8)
    status = capi.new_status()
    ukeys, nu, bounds, eb, local = capi.plan_unique_rows(d_ids, v, shards, ws, status=status, want_elem_bounds=True)
    plan = _Plan()
    wk, wn, wb, web, wl = CpuCheckerOps().plan_unique(torch.from_numpy(ids), v, shards, plan, True)
    u = int(nu.item())
    assert u == int(wn.item()) and int(status.item()) & 2
    assert np.array_equal(ukeys.cpu().numpy()[:u], wk.numpy()[:u])
    assert np.array_equal(local.cpu().numpy()[:u], wl.numpy()[:u])
    assert bounds.cpu().tolist() == wb.tolist()
    assert eb.cpu().tolist() == web.tolist()
    # owner of every unique key agrees with the reference's id % shard_num
    hs = (v + shards - 1) // shards
    if shards > 1:
        gids = local.cpu().numpy()[:u] * shards + ukeys.cpu().numpy()[:u] // hs
        assert np.array_equal(np.sort(gids), np.unique(ids[ids < v]))
        per_owner = oracle.split_ids(np.unique(ids[ids < v]), shards)
        b = bounds.cpu().numpy()
        for g in range(shards):
            assert b[g + 1] - b[g] == len(per_owner[g])
    # expand (merge_ids) and gather_sorted against numpy on the same plan
    ph = capi.plan_height(v, shards)
    uvals = torch.from_numpy(rng.standard_normal((max(u, 1), 8)).astype(np.float32)).cuda()
    got = capi.expand_unique(uvals, n, ph, ws).cpu().numpy()
    valid = ids < v
    want = uvals.cpu().numpy()[np.minimum(plan.uidx, max(u, 1) - 1)][np.argsort(plan.perm, kind="stable")]
    assert np.array_equal(got[valid], want[valid])
    assert not got[~valid].any()          # out-of-range ids come back as zero rows, never another id's row
    val = torch.from_numpy(rng.standard_normal((n, 8)).astype(np.float32)).cuda()
    assert np.array_equal(capi.gather_sorted(val, ph, ws).cpu().numpy(), val.cpu().numpy()[plan.perm])


@pytest.mark.parametrize("shards", [1, 4])
def test_expand_unique_all_ids_out_of_range(capi, shards):
    """A batch with no valid id at all: U = 0, every expanded row is zero (nothing is read at a wrapped index)."""
    n, v = 3000, 500
    ids = torch.full((n,), v + 7, dtype=torch.int64, device="cuda")
    ids[::3] = -1
    ws = capi.alloc_workspace(n, 8)
    status = capi.new_status()
    ukeys, nu, bounds, eb, local = capi.plan_unique_rows(ids, v, shards, ws, status=status, want_elem_bounds=True)
    assert int(nu.item()) == 0 and int(status.item()) & 2
    assert bounds.cpu().tolist() == [0] * (shards + 1) and eb.cpu().tolist() == [0] * (shards + 1)
    uvals = torch.full((4, 8), 3.5, device="cuda")
    got = capi.expand_unique(uvals, n, capi.plan_height(v, shards), ws)
    torch.cuda.synchronize()
    assert not got.cpu().numpy().any()


@pytest.mark.parametrize("deterministic", [True, False], ids=["bit-exact", "fast"])
def test_sharded_table_world1_cuda(capi, pkg, deterministic):
    v, d, n = 50000, 64, 30000
    rng = np.random.default_rng(5)
    full = (rng.random((v, d)) * 0.1 - 0.05).astype(np.float32)
    ids = _zipf_ids(rng, n, v)
    dout = (rng.standard_normal((n, d)) * 0.01).astype(np.float32)
    tab = pkg.sharded.ShardedTable(v, d, pkg.sharded.CudaOps(capi), torch.device("cuda"),
                                   init=lambda rows: torch.from_numpy(full[rows.numpy()]))
    ws = capi.alloc_workspace(n, d)
    ids_d = torch.from_numpy(ids).cuda()
    # the fast variant issues the ids-only part ahead of time on a side stream (the bench's prefetch)
    prep = None if deterministic else tab.prepare(ids_d, ws, deterministic, stream=torch.cuda.Stream())
    out, ctx = tab.lookup(ids_d, ws, deterministic=deterministic, prep=prep)
    assert np.array_equal(out.cpu().numpy(), full[ids])
    lr = torch.tensor([0.05], device="cuda")
    tab.update(ctx, torch.from_numpy(dout).cuda(), lr, deterministic=deterministic)
    after = tab.gather_full_table().cpu().numpy()
    want = full.copy()
    oracle.sgd_sparse(want, ids, dout, np.array([0.05], np.float32))
    if deterministic:
        assert np.array_equal(after.view(np.uint32), want.view(np.uint32))
    else:
        rows, merged = oracle.merge_add(ids, dout, truth=True)
        truth = full.astype(np.float64)
        truth[rows] -= 0.05 * merged
        _, l1 = oracle.merge_add(ids, np.abs(dout), truth=True)
        bound = 1e-6 * np.abs(full)
        bound[rows] += 1e-6 * 0.05 * l1
        assert (np.abs(after - truth) <= bound + 1e-30).all()


def _nccl_worker(rank, world, port, deterministic, q, peer=False):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    try:
        import __graft_entry__ as ge
        pkg = ge.load_package()
        v, d, n = 200003, 64, 60000
        full = (np.random.default_rng(7).random((v, d)) * 0.1 - 0.05).astype(np.float32)
        rng = np.random.default_rng(100 + rank)
        ids = _zipf_ids(rng, n, v)
        dout = (rng.standard_normal((n, d)) * 0.01).astype(np.float32)
        dev = torch.device("cuda", rank)
        tab = pkg.sharded.ShardedTable(v, d, pkg.sharded.CudaOps(pkg.capi), dev,
                                       init=lambda rows: torch.from_numpy(full[rows.numpy()]))
        ws = pkg.capi.alloc_workspace(n, d, 4, dev)
        if peer:
            assert tab.enable_engine(n, 2 * n)
        ids_d = torch.from_numpy(ids).to(dev)
        prep = tab.prepare(ids_d, ws, deterministic, stream=torch.cuda.Stream(dev)) if peer else None
        out, ctx = tab.lookup(ids_d, ws, deterministic=deterministic, prep=prep)
        ok = np.array_equal(out.cpu().numpy(), full[ids]) and (ctx["fused"] == (peer and not deterministic))
        tab.update(ctx, torch.from_numpy(dout).to(dev), torch.tensor([0.05], device=dev), deterministic=deterministic)
        after = tab.gather_full_table().cpu().numpy()
        if peer:   # no arrival flag timed out, no capacity was exceeded
            ok = ok and tab.check_status() == 0
        q.put((rank, ok, after if rank == 0 else None, ids, dout))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("deterministic,peer", [(True, False), (False, False), (False, True)],
                         ids=["bit-exact", "fast-nccl", "fast-engine"])
def test_sharded_table_two_gpus_nccl(capi, deterministic, peer):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, deterministic, q, peer)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(g[1] for g in got)
    v, d = 200003, 64
    want = (np.random.default_rng(7).random((v, d)) * 0.1 - 0.05).astype(np.float32)
    base = want.copy()
    ids = np.concatenate([g[3] for g in got])
    dout = np.concatenate([g[4] for g in got])
    after = got[0][2]
    if deterministic:
        oracle.sgd_sparse(want, ids, dout, np.array([0.05], np.float32))
        assert np.array_equal(after.view(np.uint32), want.view(np.uint32))
    else:
        rows, merged = oracle.merge_add(ids, dout, truth=True)
        truth = base.astype(np.float64)
        truth[rows] -= 0.05 * merged
        _, l1 = oracle.merge_add(ids, np.abs(dout), truth=True)
        bound = 1e-6 * np.abs(base)
        bound[rows] += 1e-6 * 0.05 * l1
        assert (np.abs(after - truth) <= bound + 1e-30).all()


def test_split_ids_merge_ids_reference_semantics(capi, goldens):
    """Device split_ids / merge_ids against the reference's own test vectors (G9: test_split_ids_op.py:26-38,
    test_merge_ids_op.py:22-34) and against the oracle on random ids."""
    g = goldens["G9_split_merge_ids"]
    ids = np.array(g["ids"], np.int64)
    ws = capi.alloc_workspace(ids.size, 2)
    out, bounds = capi.split_ids(torch.from_numpy(ids).cuda(), 100, g["shard_num"], ws)
    b = bounds.cpu().tolist()
    got = [out.cpu().numpy()[b[s]:b[s + 1]].tolist() for s in range(g["shard_num"])]
    assert got == g["expect_shards"]
    x = np.concatenate([np.array(x_, np.float32).reshape(-1, 2) for x_ in g["merge_x"]])
    merged = capi.merge_ids(torch.from_numpy(x).cuda(), g["shard_num"], ws).cpu().numpy()
    assert np.array_equal(merged, np.array(g["merge_expect"], np.float32))
    rng = np.random.default_rng(9)
    for n, v, shards, d in [(5000, 977, 4, 16), (200000, 1 << 20, 8, 64), (3000, 50, 3, 5)]:
        ids = _zipf_ids(rng, n, v)
        ws = capi.alloc_workspace(n, d)
        out, bounds = capi.split_ids(torch.from_numpy(ids).cuda(), v, shards, ws)
        want = oracle.split_ids(ids, shards)
        b = bounds.cpu().tolist()
        for s in range(shards):
            assert np.array_equal(out.cpu().numpy()[b[s]:b[s + 1]], want[s])
        blocks = [rng.standard_normal((len(w), d)).astype(np.float32) for w in want]
        x = np.concatenate(blocks)
        merged = capi.merge_ids(torch.from_numpy(x).cuda(), shards, ws).cpu().numpy()
        assert np.array_equal(merged, oracle.merge_ids(ids, blocks))


def test_exchange_maps_matches_numpy(capi):
    """b200emb_exchange_maps is a pure function of the gathered split table: check it against numpy for
    every rank of a made-up 5-rank table (the pointers are just integers here)."""
    rng = np.random.default_rng(9)
    world, row_bytes = 5, 256
    cols = 2 * (world + 1)
    tab = np.zeros((world, cols), np.int64)
    tab[:, 1:world + 1] = np.cumsum(rng.integers(0, 1000, (world, world)), axis=1)
    pi, pu, pg = (rng.integers(1 << 20, 1 << 40, world) * 256 for _ in range(3))
    tab_d = torch.from_numpy(tab).cuda()
    ptrs = [torch.from_numpy(np.asarray(p, np.int64)).cuda() for p in (pi, pu, pg)]
    for rank in range(world):
        got = capi.exchange_maps(tab_d, rank, ptrs[0], ptrs[1], ptrs[2], row_bytes).cpu().numpy()
        before = (tab[:rank, 1:world + 1] - tab[:rank, :world]).sum(axis=0)
        cnt = tab[:, rank + 1] - tab[:, rank]
        assert np.array_equal(got[0, :world], pi + 8 * before)
        assert np.array_equal(got[1], np.concatenate([[0], np.cumsum(cnt)]))
        assert np.array_equal(got[2, :world], pu + row_bytes * tab[:, rank])
        assert np.array_equal(got[3, :world], pg + row_bytes * before)


def test_peer_barrier_and_allgather_single_rank(capi):
    """world = 1: the peer kernels talk to this GPU's own buffers (the multi-GPU tests cover the real thing)."""
    flags = torch.zeros(32, dtype=torch.int64, device="cuda")
    table = torch.zeros((1, 6), dtype=torch.int64, device="cuda")
    fp = torch.tensor([flags.data_ptr()], dtype=torch.int64, device="cuda")
    tp = torch.tensor([table.data_ptr()], dtype=torch.int64, device="cuda")
    status = capi.new_status()
    for epoch in (1, 2, 3):
        capi.peer_barrier(fp, 1, 0, epoch, status)
    assert int(flags[0].item()) == 3
    row = torch.arange(10, 16, dtype=torch.int64, device="cuda")
    host = torch.zeros((1, 6), dtype=torch.int64).pin_memory()
    capi.peer_allgather_i64(row, tp, table, fp, 1, 0, 4, host_copy=host, status=status)
    torch.cuda.synchronize()
    assert host.reshape(-1).tolist() == list(range(10, 16)) and torch.equal(table.reshape(-1), row)
    assert int(status.item()) == 0
    with pytest.raises(capi.B200EmbError):
        capi.peer_barrier(fp, 1, 1, 5, status)       # rank outside the world
