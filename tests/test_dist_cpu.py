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
