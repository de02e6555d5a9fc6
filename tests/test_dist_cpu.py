"""world_size-2 gloo test (CPU) of the data-parallel exchange in deep-ctr_b200/dist.py: after the sparse
all-gather + dense all-reduce, every rank holds exactly what a single process would compute on the
concatenated (rank-major) global batch -- checked with the oracle's own update rules."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    import deepctr_oracle as O
    from deepctr_b200 import dist as ddist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(7)               # same stream on every rank: the GLOBAL batch
        X, F, d, N = 40, 4, 3, 6
        table = rng.normal(0, 0.5, (X, d))
        ids = rng.integers(0, 10, (world * N, F))
        ge = rng.normal(0, 1, (world * N, F, d))
        dense = rng.normal(0, 1, (world, 17))
        lo, hi = rank * N, (rank + 1) * N
        gids, gwts, gg = ddist.exchange_sparse(dist, torch.as_tensor(ids[lo:hi]), None, torch.as_tensor(ge[lo:hi]))
        assert gwts is None
        assert np.array_equal(gids.numpy(), ids) and np.array_equal(gg.numpy(), ge)      # rank-major, bit-exact
        flat = ddist.allreduce_dense(dist, torch.as_tensor(dense[rank].copy()))
        np.testing.assert_allclose(flat.numpy(), dense.sum(0), rtol=1e-12)
        # the update every rank applies == single-process update on the global batch (both semantics)
        a, b = table.copy(), table.copy()
        O.seq_decay_scatter(a, gids.numpy(), gg.numpy(), 0.1, 0.99)
        O.seq_decay_scatter(b, ids, ge, 0.1, 0.99)
        assert np.array_equal(a, b)
        tabs = ddist.all_gather_cat(dist, torch.as_tensor(a[None]))
        assert np.array_equal(tabs[0].numpy(), tabs[1].numpy())                          # replicas stay identical
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_dp_exchange_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def _shard_worker(rank, world, port, q):
    """row-sharded lookup plumbing (plan_exchange + the three all-to-alls) with numpy standing in for the kernels"""
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from deepctr_b200 import dist as ddist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(11)
        X, d, N, F = 37, 3, 5, 4                                  # X not a multiple of world: the last shard is short
        table = rng.normal(0, 1, (X, d))
        ids_all = rng.integers(0, X, (world, N * F))
        ids_all[1, :3] = ids_all[0, :3]                           # rows requested by both ranks
        mine = torch.as_tensor(table[rank::world].copy())
        flat = torch.as_tensor(ids_all[rank])
        perm, counts, send_rows = ddist.plan_exchange(flat, world, lambda k: torch.argsort(k, stable=True))
        owner = ids_all[rank] % world
        assert np.array_equal(counts.numpy(), np.bincount(owner, minlength=world))
        assert np.array_equal(owner[perm.numpy()], np.sort(owner))                      # bucketed by owner ...
        for r in range(world):
            assert np.all(np.diff(perm.numpy()[np.sort(owner) == r]) > 0)               # ... slot order kept inside
        recv_counts = ddist.exchange_counts(dist, counts)
        sc, rc = counts.tolist(), recv_counts.tolist()
        assert rc == [int(np.sum(ids_all[s] % world == rank)) for s in range(world)]
        recv_rows = ddist.all_to_all_rows(dist, send_rows, sc, rc)
        want = np.concatenate([ids_all[s][ids_all[s] % world == rank] // world for s in range(world)])
        assert np.array_equal(recv_rows.numpy(), want)                                  # (source rank, slot) order
        back = ddist.all_to_all_rows(dist, mine[recv_rows], rc, sc)
        inv = torch.empty(N * F, dtype=torch.long)
        inv[perm] = torch.arange(N * F)
        assert np.array_equal(back[inv].numpy(), table[ids_all[rank]])                  # every slot got its global row
        # gradient route: per-slot values travel to the owner aligned with recv_rows
        g = torch.as_tensor(rng.normal(0, 1, (world, N * F, d))[rank])
        recv_g = ddist.all_to_all_rows(dist, g[perm], sc, rc)
        acc = np.zeros_like(table[rank::world])
        np.add.at(acc, recv_rows.numpy(), recv_g.numpy())
        rng2 = np.random.default_rng(11)
        rng2.normal(0, 1, (X, d)); rng2.integers(0, X, (world, N * F))
        g_all = rng2.normal(0, 1, (world, N * F, d))
        full = np.zeros_like(table)
        np.add.at(full, ids_all.reshape(-1), g_all.reshape(-1, d))
        np.testing.assert_allclose(acc, full[rank::world], rtol=1e-12, atol=1e-12)
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_sharded_exchange_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
