"""GPU tests of the row-sharded step engine (b200emb_shard_*, csrc/shard_engine.cu) through the C ABI.

The engine replaces ReduceOpHandle / BroadcastOpHandle for sparse gradients
(framework/details/reduce_op_handle.cc:88-94, broadcast_op_handle.cc:71-122); its result is checked against
the oracle on the UNSHARDED table, the way reduce_op_handle_test.cc:153-209 checks the gathered SelectedRows
against the concatenation of every device's rows.

* world 1/2/4 with all ranks hosted on ONE GPU: every phase of the step is issued for all ranks before the
  next phase (kernels that wait for each other must not share a GPU out of order), several consecutive steps,
  rotating slots, prepare() of the next batch issued early on a side stream.  Runs on a 1-GPU box.
* world 2 (and 4/8 when the box has them) with one process per GPU and CUDA-IPC blocks: full lookup()/update(),
  one rank artificially slowed down to shake out cross-step races.
"""
import os
import socket
import sys

import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def capi(pkg):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg.capi.lib()
    return pkg.capi


def _zipf_ids(rng, n, v, a=1.1):
    w = np.arange(1, v + 1, dtype=np.float64) ** -a
    cdf = np.cumsum(w) / w.sum()
    return ((np.searchsorted(cdf, rng.random(n)) * 2654435761 + 12345) % v).astype(np.int64)


def _shard_of(full, rank, world):
    hs = (full.shape[0] + world - 1) // world
    out = np.zeros((hs, full.shape[1]), np.float32)
    rows = np.arange(rank, full.shape[0], world)
    out[: len(rows)] = full[rows]
    return out


def _unshard(shards, height):
    world = len(shards)
    full = np.stack(shards, axis=1).reshape(world * shards[0].shape[0], -1)
    return full[:height]


def _fast_sgd_bound(base, ids, dout, lr):
    """fp64 evaluation of the step and the reordering bound of the fast mode."""
    rows, merged = oracle.merge_add(ids, dout, truth=True)
    truth = base.astype(np.float64)
    truth[rows] -= lr * merged
    _, l1 = oracle.merge_add(ids, np.abs(dout), truth=True)
    bound = 1e-6 * np.abs(base).astype(np.float64)
    bound[rows] += 1e-6 * lr * l1
    return truth, bound


@pytest.mark.parametrize("world,optimizer", [(1, "sgd"), (2, "sgd"), (4, "sgd"), (3, "adam")])
def test_engine_ranks_on_one_gpu(capi, world, optimizer):
    v, d, n, steps = 100003, 64, 20000, 4
    rng = np.random.default_rng(world * 7 + 1)
    full = (rng.random((v, d)) * 0.1 - 0.05).astype(np.float32)
    dev = torch.device("cuda")
    shards = [torch.from_numpy(_shard_of(full, r, world)).to(dev) for r in range(world)]
    m1 = [torch.zeros_like(s) for s in shards] if optimizer == "adam" else [None] * world
    m2 = [torch.zeros_like(s) for s in shards] if optimizer == "adam" else [None] * world
    eng = [capi.ShardEngine(world, r, v, d, shards[r], n, 2 * n, optimizer, m1[r], m2[r]) for r in range(world)]
    capi.ShardEngine.connect_local(eng)
    lr = torch.tensor([0.05], device=dev)
    b1p, b2p = torch.tensor([0.9], device=dev), torch.tensor([0.999], device=dev)
    side = torch.cuda.Stream(dev)
    ref = full.copy()
    rm1, rm2 = np.zeros_like(full), np.zeros_like(full)
    batches = [[_zipf_ids(rng, n - 37 * r, v) for r in range(world)] for _ in range(steps)]
    for b in batches:               # a few out-of-range ids: zero rows out, no update, status bit 1
        b[0][::501] = v + 3
    grads = [[(rng.standard_normal((len(i), d)) * 0.01).astype(np.float32) for i in b] for b in batches]
    ids_d = [[torch.from_numpy(i).to(dev) for i in b] for b in batches]
    g_d = [[torch.from_numpy(g).to(dev) for g in gs] for gs in grads]

    main = torch.cuda.current_stream(dev)

    def prepare(step):
        side.wait_stream(main)
        with torch.cuda.stream(side):
            for r in range(world):
                eng[r].prepare(step % 4, ids_d[step][r])

    prepare(0)
    for step in range(steps):
        slot = step % 4
        torch.cuda.current_stream(dev).wait_stream(side)
        if step + 1 < steps:
            prepare(step + 1)           # ids-only work of the next batch, beside this step
        outs = [torch.empty((len(batches[step][r]), d), device=dev) for r in range(world)]
        for ph in range(4):
            for r in range(world):
                eng[r].lookup(slot, outs[r], phases=(ph, ph))
        for r in range(world):
            ids = batches[step][r]
            want = ref[np.clip(ids, 0, v - 1)]
            want[(ids < 0) | (ids >= v)] = 0
            assert np.array_equal(outs[r].cpu().numpy().view(np.uint32), want.view(np.uint32)), (step, r)
        for ph in range(2):
            for r in range(world):
                eng[r].update(slot, g_d[step][r], lr, b1p, b2p, phases=(ph, ph))
        torch.cuda.synchronize()
        all_ids = np.concatenate(batches[step])
        all_g = np.concatenate(grads[step])
        ok = (all_ids >= 0) & (all_ids < v)
        after = _unshard([s.cpu().numpy() for s in shards], v)
        if optimizer == "sgd":
            truth, bound = _fast_sgd_bound(ref, all_ids[ok], all_g[ok], 0.05)
            assert (np.abs(after - truth) <= bound + 1e-30).all(), step
        else:
            wp, w1, w2 = ref.copy(), rm1.copy(), rm2.copy()
            oracle.adam_sparse(wp, w1, w2, all_ids[ok], all_g[ok], np.array([0.05], np.float32),
                               np.array([0.9], np.float32), np.array([0.999], np.float32))
            # Adam divides by sqrt(m2): where a merged gradient nearly cancels, the rounding of the fast-mode
            # tree moves the parameter by a visible fraction of one step (lr_t * m1 / sqrt(m2) is O(lr)), so the
            # parameters are held to a small fraction of outliers and the moments (linear in g) to the bound
            diff = np.abs(after - wp)
            assert (diff > 2e-5 * np.abs(wp) + 1e-7).mean() < 1e-5 and diff.max() < 0.1
            rm1 = _unshard([m.cpu().numpy() for m in m1], v)
            rm2 = _unshard([m.cpu().numpy() for m in m2], v)
            _, l1 = oracle.merge_add(all_ids[ok], np.abs(all_g[ok]), truth=True)
            rows_u = np.unique(all_ids[ok])
            bound = 1e-6 * np.abs(w1)
            bound[rows_u] += 1e-6 * 0.1 * l1
            assert (np.abs(rm1 - w1) <= bound + 1e-30).all()
        ref = after          # the next step starts from what the engine left (errors do not accumulate in the check)
    for r in range(world):
        st = int(eng[r].status.item())
        assert st & capi.STATUS_BARRIER_TIMEOUT == 0 and st & capi.STATUS_SHARD_OVERFLOW == 0
        assert (st & 2) == (2 if r == 0 else 0)      # rank 0's out-of-range ids were flagged by its plan
        bounds, nu, seg, epoch = eng[r].debug_views((steps - 1) % 4)
        assert int(epoch.item()) == 1 and int(bounds[-1].item()) == int(nu.item())   # every slot was used once
    for e in eng:
        e.destroy()


def test_engine_capacity_overflow_is_flagged(capi):
    """More unique rows than max_recv_rows: the excess is dropped and the status word says so (no fault)."""
    v, d, n = 5000, 16, 4000
    dev = torch.device("cuda")
    shard = torch.zeros((v, d), device=dev)
    e = capi.ShardEngine(1, 0, v, d, shard, n, 100)
    capi.ShardEngine.connect_local([e])
    ids = torch.arange(n, dtype=torch.int64, device=dev)
    e.prepare(0, ids)
    e.lookup(0)
    e.update(0, torch.ones((n, d), device=dev), torch.tensor([1.0], device=dev))
    torch.cuda.synchronize()
    assert int(e.status.item()) & capi.STATUS_SHARD_OVERFLOW
    with pytest.raises(capi.B200EmbError):
        e.check_status()
    e.destroy()


# ------------------------------------------------------------------ one process per GPU
def _worker(rank, world, port, q, steps, n, slow_rank):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)     # the handle exchange needs no NCCL
    try:
        import __graft_entry__ as ge
        capi = ge.load_package().capi
        v, d = 200003, 64
        full = (np.random.default_rng(7).random((v, d)) * 0.1 - 0.05).astype(np.float32)
        dev = torch.device("cuda", rank)
        shard = torch.from_numpy(_shard_of(full, rank, world)).to(dev)
        eng = capi.ShardEngine(world, rank, v, d, shard, n, 2 * n)
        handles = [None] * world
        dist.all_gather_object(handles, eng.handle)
        eng.connect(handles)
        rng = np.random.default_rng(100 + rank)
        lr = torch.tensor([0.05], device=dev)
        side = torch.cuda.Stream(dev)
        ids_np = [_zipf_ids(rng, n, v) for _ in range(steps)]
        g_np = [(rng.standard_normal((n, d)) * 0.01).astype(np.float32) for _ in range(steps)]
        ids_d = [torch.from_numpy(i).to(dev) for i in ids_np]
        g_d = [torch.from_numpy(g).to(dev) for g in g_np]
        outs = []
        with torch.cuda.stream(side):
            eng.prepare(0, ids_d[0])
        for s in range(steps):
            torch.cuda.current_stream(dev).wait_stream(side)
            if s + 1 < steps:
                side.wait_stream(torch.cuda.current_stream(dev))
                with torch.cuda.stream(side):
                    eng.prepare((s + 1) % 4, ids_d[s + 1])
            if slow_rank is not None and rank == slow_rank:
                torch.cuda._sleep(int(3e6))          # ~1.5 ms: this rank lags, the others run ahead
            outs.append(eng.lookup(s % 4))
            if slow_rank is not None and rank == (slow_rank + 1) % world:
                torch.cuda._sleep(int(2e6))
            eng.update(s % 4, g_d[s], lr)
        torch.cuda.synchronize()
        dist.barrier()
        st = int(eng.status.item())
        q.put((rank, st, shard.cpu().numpy(), ids_np, g_np, [o.cpu().numpy() for o in outs]))
        dist.barrier()
        eng.destroy()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("slow_rank", [None, 1], ids=["even", "skewed"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_engine_one_process_per_gpu(capi, world, slow_rank):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (run with gpurun --gpus %d)" % (world, world))
    import torch.multiprocessing as mp
    steps, n = 8, 30000
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, steps, n, slow_rank)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=600) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(g[1] == 0 for g in got), [g[1] for g in got]
    v, d = 200003, 64
    ref = (np.random.default_rng(7).random((v, d)) * 0.1 - 0.05).astype(np.float32)
    bound_acc = np.zeros((v, d))
    truth = ref.astype(np.float64)
    for s_ in range(steps):
        ids = np.concatenate([g[3][s_] for g in got])
        dout = np.concatenate([g[4][s_] for g in got])
        # forward of step s_ saw the table after s_ updates: within the accumulated reordering bound of the truth
        for g in got:
            o, i_ = g[5][s_], g[3][s_]
            bad = np.abs(o - truth[i_]) > bound_acc[i_] + 1e-6 * np.abs(truth[i_]) + 1e-30
            if bad.any():     # say what the rows look like instead: another step's ids? the untouched table?
                alt = {}
                for t_ in range(steps):
                    alt["ids_of_step_%d" % t_] = float((np.abs(o - truth[g[3][t_]]) < 1e-3).all(axis=1).mean())
                alt["initial_table_same_ids"] = float((o == ref[i_]).all(axis=1).mean())
                raise AssertionError("step %d rank %d: %d of %d elements off (max %.3g); rows owned by ranks %s; "
                                     "fraction of rows matching %s" % (
                                         s_, g[0], int(bad.sum()), bad.size, float(np.abs(o - truth[i_]).max()),
                                         sorted(set((i_[bad.any(axis=1)] % world).tolist())), alt))
        rows, merged = oracle.merge_add(ids, dout, truth=True)
        _, l1 = oracle.merge_add(ids, np.abs(dout), truth=True)
        truth[rows] -= 0.05 * merged
        bound_acc[rows] += 1e-6 * 0.05 * l1
    after = _unshard([g[2] for g in got], v)
    assert (np.abs(after - truth) <= bound_acc + 1e-6 * np.abs(truth) + 1e-30).all()
