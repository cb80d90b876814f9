"""CPU checker backend for paddle-1_b200/sharded.py — TEST INFRASTRUCTURE ONLY.

Implements the `ops` interface of ShardedTable with numpy + the oracle so that the exchange logic
(split sizes, send order, merge_ids semantics, bit-exact ordering) can be exercised with gloo and
world_size 2 on a machine without GPUs.  The product never imports this module."""
import numpy as np
import torch

import oracle


class _Plan:
    pass


class CpuCheckerOps:
    def alloc_workspace(self, n, d, elem_size, device):
        return _Plan()

    def plan_unique(self, ids, height, world, ws, want_elem_bounds):
        i = ids.numpy().astype(np.int64)
        hs = (height + world - 1) // world
        key = (i % world) * hs + i // world if world > 1 else i.copy()
        plan_h = world * hs if world > 1 else height
        key = np.where((i >= 0) & (i < height), key, plan_h)
        perm = np.argsort(key, kind="stable")
        skey = key[perm]
        valid = skey < plan_h
        ukeys, first = np.unique(skey[valid], return_index=True)
        uidx = np.searchsorted(ukeys, skey)            # unique index of every sorted position
        ws.perm, ws.uidx, ws.skey, ws.key = perm, uidx, skey, key
        n = i.size
        out_keys = np.zeros(max(n, 1), np.int64)
        out_keys[: ukeys.size] = ukeys
        local = np.zeros(max(n, 1), np.int64)
        local[: ukeys.size] = ukeys % (hs if world > 1 else plan_h)
        targets = np.arange(world + 1, dtype=np.int64) * (hs if world > 1 else plan_h)
        bounds = np.searchsorted(ukeys, targets)
        bounds[world] = ukeys.size
        eb = None
        if want_elem_bounds:
            eb = np.searchsorted(skey, targets)
            eb[world] = int(valid.sum())
            eb = torch.from_numpy(eb.astype(np.int64))
        return (torch.from_numpy(out_keys), torch.tensor([ukeys.size]), torch.from_numpy(bounds.astype(np.int64)),
                eb, torch.from_numpy(local))

    def expand_unique(self, unique_values, n, plan_h, ws):
        out = np.empty((n, unique_values.shape[1]), unique_values.numpy().dtype)
        out[ws.perm] = unique_values.numpy()[ws.uidx]
        return torch.from_numpy(out)

    def gather_sorted(self, value, plan_h, ws):
        return torch.from_numpy(value.numpy()[ws.perm].copy())

    def merge_by_plan(self, ids, value, plan_h, ws):
        rows, merged = oracle.merge_add(ws.key[ws.key < plan_h], value.numpy()[ws.key < plan_h])
        out = np.zeros((ids.numel(), value.shape[1]), value.numpy().dtype)
        out[: rows.size] = merged
        return torch.from_numpy(out)

    def lookup(self, shard, local_rows):
        return torch.from_numpy(oracle.lookup_table(shard.numpy(), local_rows.numpy()))

    def sgd(self, shard, rows, grads, lr, deterministic, ws=None):
        a = shard.numpy()
        if deterministic:
            oracle.sgd_sparse(a, rows.numpy(), grads.numpy(), lr.numpy())
        else:  # merged update, fp64 accumulate (what the fast mode is held to)
            r, m = oracle.merge_add(rows.numpy(), grads.numpy(), truth=True)
            a[r] = (a[r].astype(np.float64) - float(lr.numpy()[0]) * m).astype(a.dtype)
