"""torch.ops.deepctr.* (deep-ctr_b200/ops.py): schema checks (torch.library.opcheck: the declared mutations are the real ones),
boundary validation (dtype / device / layout / alignment) and equality with the raw C-ABI call."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _emb_case():
    from deepctr_b200 import _lib as L, ops  # noqa: F401
    g = torch.Generator(device="cuda").manual_seed(1)
    X, F, d, ld, N = 500, 6, 5, 8, 37
    table = torch.zeros(X, ld, device="cuda")
    table[:, :d].uniform_(-1, 1, generator=g)
    ids = torch.randint(0, X, (N, F), device="cuda", dtype=torch.int32, generator=g)
    ids[3] = ids[2]
    wts = torch.rand(N, F, device="cuda", generator=g)
    bias = torch.tensor([0.3], device="cuda")
    return X, F, d, ld, N, table, ids, wts, bias


def test_opcheck_schemas():
    from deepctr_b200 import _lib as L
    from torch.library import opcheck
    O = torch.ops.deepctr
    X, F, d, ld, N, table, ids, wts, bias = _emb_case()
    D0 = F * d + 1
    out = torch.zeros(N, 32, device="cuda")
    opcheck(O.gather_concat_fwd.default, (ids, wts, table, d, 0, False, bias, 1, out), test_utils=("test_schema",))
    M = N * F
    rows, pos = torch.zeros(M, dtype=torch.int32, device="cuda"), torch.zeros(M, dtype=torch.int32, device="cuda")
    ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    opcheck(O.sort_pairs.default, (ids, None, M, X, rows, pos, ws), test_utils=("test_schema",))
    O.sort_pairs(ids, None, M, X, rows, pos, ws)
    fws = torch.zeros(L.lib().dctr_scatter_workspace(M, ld), dtype=torch.uint8, device="cuda")
    g = torch.randn(N, 32, device="cuda")
    opcheck(O.scatter_sgd.default, (0, rows, pos, M, F, d, g, d, wts, -1, 0.1, 1.0, table.clone(), fws), test_utils=("test_schema",))
    p, gr = torch.ones(64, device="cuda"), torch.ones(64, device="cuda")
    opcheck(O.dense_sgd.default, (p, gr, 64, 0.5, 1.0), test_utils=("test_schema",))
    opcheck(O.vec_sum.default, (gr, 64, 1.0, torch.zeros(1, device="cuda")), test_utils=("test_schema",))


def test_ops_equal_raw_abi_and_validate():
    from deepctr_b200 import _lib as L
    O = torch.ops.deepctr
    X, F, d, ld, N, table, ids, wts, bias = _emb_case()
    a, b = torch.zeros(N, 32, device="cuda"), torch.zeros(N, 32, device="cuda")
    O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, a)
    L.call("dctr_gather_concat_fwd", ids.data_ptr(), wts.data_ptr(), table.data_ptr(), X, ld, N, F, d, 0, 0, bias.data_ptr(), 1,
           b.data_ptr(), 32, None)
    torch.cuda.synchronize()
    assert torch.equal(a, b) and a.abs().sum() > 0
    # a column-offset view carries its own pointer and leading dimension
    big = torch.zeros(N, 40, device="cuda")
    O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, big[:, 4:36])
    assert torch.equal(big[:, 4:36], a) and not big[:, :4].any() and not big[:, 36:].any()
    with pytest.raises(TypeError):
        O.gather_concat_fwd(ids.long(), wts, table, d, 0, False, bias, 1, a)
    with pytest.raises(ValueError):
        O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, big[:, 1:33])        # not 16-byte aligned
    with pytest.raises(ValueError):
        O.gather_concat_fwd(ids.t().contiguous().t(), wts, table, d, 0, False, bias, 1, a)    # innermost dimension not contiguous
    with pytest.raises(NotImplementedError):
        O.gather_concat_fwd(ids.cpu(), wts.cpu(), table.cpu(), d, 0, False, bias.cpu(), 1, a.cpu())
    with pytest.raises(RuntimeError, match="gather_concat_fwd"):                        # the C-ABI's own EINVAL surfaces as RuntimeError
        O.gather_concat_fwd(ids, wts, table, d, 7, False, bias, 1, a)
    # the op launches on the CURRENT stream of the tensors' device
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        c = torch.zeros(N, 32, device="cuda")
        O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, c)
    s.synchronize()
    assert torch.equal(c, a)


def test_dropout_masks_kernel_and_graph_replay():
    """dctr_dropout_masks: values in {0, 1/keep}, keep fraction, fresh masks on every launch AND every graph replay (the step
    counter lives on the device), reproducible for the same (seed, step), first block applied in place."""
    O = torch.ops.deepctr
    n, keep = 1 << 20, 0.5
    m = torch.zeros(n, device="cuda")
    st = torch.zeros(2, dtype=torch.int32, device="cuda")
    x = torch.ones(4096, device="cuda")
    O.dropout_masks(m, keep, 7, st, x)
    torch.cuda.synchronize()
    first = m.clone()
    assert set(torch.unique(m).tolist()) == {0.0, 2.0}
    assert abs(float((m > 0).float().mean()) - keep) < 3e-3
    assert torch.equal(x, m[:4096]) and int(st[0]) == 1 and int(st[1]) == 0
    O.dropout_masks(m, keep, 7, st, None)
    assert not torch.equal(m, first) and int(st[0]) == 2
    st.zero_()
    O.dropout_masks(m, keep, 7, st, None)
    assert torch.equal(m, first)                      # same (seed, step) -> same masks
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g):
            O.dropout_masks(m, 0.8, 7, st, None)
    seen = []
    for _ in range(3):
        g.replay()
        torch.cuda.synchronize()
        seen.append(m.clone())
    assert not torch.equal(seen[0], seen[1]) and not torch.equal(seen[1], seen[2])
    assert abs(float((seen[2] > 0).float().mean()) - 0.8) < 3e-3 and abs(float(seen[2].max()) - 1.25) < 1e-6


def test_tf_dropout_runs_inside_the_captured_step(tmp_path):
    """FNN with keep_prob 0.5 (baseline.py:164): the masks are drawn by the library inside the CUDA graph -- no eager torch op per
    step -- and differ from step to step; predictions use no dropout."""
    import deepctr_b200.models as M
    from deepctr_b200 import _lib as L
    rng = np.random.default_rng(0)
    cat_sizes = np.array([50, 7, 200])
    offsets = [2 + int(cat_sizes[:i].sum()) for i in range(3)]
    N = 256
    m = M.FNN(cat_sizes, offsets, N, [2 + int(cat_sizes.sum()), 5, 4, 32, 16, "tanh"], ["uniform", -0.1, 0.1, list(range(1, 12)), None],
              ["sgd", 0.01, "sum"], [0.5], "train", N)
    assert m.engine.dropout is not None
    feed = {m.v_wt_hldr: rng.random((N, 2)).astype(np.float32), m.c_id_hldr: np.stack([rng.integers(0, s, N) for s in cat_sizes], 1),
            m.c_wt_hldr: np.ones((N, 3), np.float32), m.lbl_hldr: rng.integers(0, 2, N)}
    masks, launches = [], []
    for step in range(5):
        l0 = L.launch_count()
        loss, p = m.train_step(feed)
        launches.append(L.launch_count() - l0)
        assert np.isfinite(loss)
        masks.append(m.engine.masks[1].clone())
    assert m._graph is not None and launches[-1] == 0          # replayed: nothing is launched from the host, masks included
    assert not torch.equal(masks[3], masks[4])
    assert abs(float((masks[4] > 0).float().mean()) - 0.5) < 0.05
    p1 = m.eval_preds.eval({m.eval_id_hldr: np.concatenate([np.broadcast_to(np.arange(2), (N, 2)), feed[m.c_id_hldr] + np.asarray(offsets)], 1),
                            m.eval_wts_hldr: np.concatenate([feed[m.v_wt_hldr], feed[m.c_wt_hldr]], 1)})
    p2 = m.eval_preds.eval({m.eval_id_hldr: np.concatenate([np.broadcast_to(np.arange(2), (N, 2)), feed[m.c_id_hldr] + np.asarray(offsets)], 1),
                            m.eval_wts_hldr: np.concatenate([feed[m.v_wt_hldr], feed[m.c_wt_hldr]], 1)})
    assert np.array_equal(p1, p2)


def test_no_writes_outside_the_buffers():
    """compute-sanitizer is closed on this pool (profiles/r02_compute_sanitizer_closed.txt), so the kernels with hand-rolled
    look-back / tickets / carries are run at odd sizes with CANARY regions on both sides of everything they write: sorted rows /
    positions, the table around the touched rows, the gather output, the fold workspace."""
    from deepctr_b200 import _lib as L
    O = torch.ops.deepctr
    g = torch.Generator(device="cuda").manual_seed(3)
    G = 64                                                   # guard elements on each side
    for M, X in [(1, 7), (2047, 100), (2049, 70001), (70001, 1 << 20), (300001, 1 << 27)]:
        F, d, ld = 1, 5, 8
        ids = torch.randint(0, X, (M,), device="cuda", dtype=torch.int32, generator=g)
        ids[: M // 3] = ids[0]                               # one long segment
        rows_g = torch.full((M + 2 * G + 8,), -7, dtype=torch.int32, device="cuda")
        pos_g = torch.full((M + 2 * G + 8,), -7, dtype=torch.int32, device="cuda")
        off = G + (4 - G % 4) % 4                            # keep the payload 16-byte aligned
        rows, pos = rows_g[off:off + M], pos_g[off:off + M]
        ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M) + 2 * G, dtype=torch.uint8, device="cuda")
        ws[-G:] = 0x5A
        O.sort_pairs(ids, None, M, X, rows, pos, ws[:-G])
        torch.cuda.synchronize()
        assert (rows_g[:off] == -7).all() and (rows_g[off + M:] == -7).all() and (pos_g[:off] == -7).all() and (pos_g[off + M:] == -7).all()
        assert (ws[-G:] == 0x5A).all()
        order = torch.argsort(ids.long(), stable=True)
        assert torch.equal(rows.long(), ids.long()[order]) and torch.equal(pos.long(), order)
        # fold into a table that only holds rows [lo, hi) of the id range, with guard rows around it
        lo, hi = int(ids.min()), int(ids.max()) + 1
        Xt = hi - lo
        if Xt * ld * 4 > (1 << 30):
            continue
        tab_g = torch.full((Xt + 2, ld), 3.0, device="cuda")
        table = tab_g[1:1 + Xt]
        grad = torch.randn(M, ld, device="cuda", generator=g)
        fws = torch.zeros(L.lib().dctr_scatter_workspace(M, ld) + G, dtype=torch.uint8, device="cuda")
        fws[-G:] = 0x5A
        local = (rows - lo).contiguous() if off % 4 == 0 else (rows - lo).clone()
        O.scatter_sgd(0, local, pos.clone(), M, 1, d, grad, ld, None, -1, 0.5, 1.0, table, fws[:-G])
        torch.cuda.synchronize()
        assert (tab_g[0] == 3.0).all() and (tab_g[-1] == 3.0).all() and (fws[-G:] == 0x5A).all()
        want = torch.full((Xt, d), 3.0, dtype=torch.float64, device="cuda")
        want.index_add_(0, (ids.long() - lo), -0.5 * grad[:, :d].double())
        assert torch.allclose(table[:, :d].double(), want, rtol=1e-5, atol=1e-5 * max(1.0, M // 3) ** 0.5)
        assert (table[:, d:] == 3.0).all()                    # pad columns of a row are not written


def test_shard_pack_buckets():
    """dctr_shard_pack: slots ordered by owner go into equal-sized buckets; inv / src_of_dest are inverse to each other, unused
    entries keep -1 / 0, an overflowing bucket is counted and its slots still point at a valid entry."""
    import deepctr_b200.ops  # noqa: F401
    O = torch.ops.deepctr
    from deepctr_b200 import _lib as L
    g = torch.Generator(device="cuda").manual_seed(21)
    for world, M, C in ((2, 1000, 600), (3, 4097, 1500), (8, 50000, 7000), (4, 777, 100)):
        ids = torch.randint(0, 100000, (M,), device="cuda", dtype=torch.int32, generator=g)
        ids[: M // 5] = ids[0]                                 # a hot row: one owner gets a fuller bucket
        owner = torch.remainder(ids, world).to(torch.int32)
        owner_sorted, perm = torch.empty_like(ids), torch.empty_like(ids)
        ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
        O.sort_pairs(owner, None, M, world, owner_sorted, perm, ws)
        send_ids = torch.full((world * C,), 7, dtype=torch.int32, device="cuda")
        src = torch.full((world * C,), 7, dtype=torch.int32, device="cuda")
        inv = torch.full((M,), -5, dtype=torch.int32, device="cuda")
        status = torch.zeros(2, dtype=torch.int32, device="cuda")
        O.shard_pack(ids, perm, owner_sorted, world, C, send_ids, src, inv, status)
        torch.cuda.synchronize()
        counts = torch.bincount(owner.long(), minlength=world)
        lost = int((counts - C).clamp_min(0).sum())
        assert int(status[0]) == lost and int(status[1]) == int(counts.max())
        assert (inv >= 0).all() and (inv < world * C).all()
        for o in range(world):
            n = min(int(counts[o]), C)
            slots = torch.nonzero(owner == o).flatten()[:n]                       # stable: original slot order inside a bucket
            assert torch.equal(src[o * C:o * C + n].long(), slots)
            assert torch.equal(send_ids[o * C:o * C + n], torch.div(ids[slots], world, rounding_mode="floor"))
            assert torch.equal(inv[slots].long(), torch.arange(o * C, o * C + n, device="cuda"))
            assert (send_ids[o * C + n:(o + 1) * C] == -1).all() and (src[o * C + n:(o + 1) * C] == 0).all()
        # the same plan with the id buckets delivered into (here: local stand-ins for) the owners' request lists
        rank = world - 1
        lists = [torch.full((world * C,), 3, dtype=torch.int32, device="cuda") for _ in range(world)]
        peers = torch.tensor([t.data_ptr() for t in lists], dtype=torch.int64, device="cuda")
        s2, r2, i2 = torch.full_like(send_ids, 7), torch.full_like(src, 7), torch.full_like(inv, -5)
        status.zero_()
        O.shard_pack_push(ids, perm, owner_sorted, world, C, s2, r2, i2, status, rank, peers, lists[rank])
        torch.cuda.synchronize()
        assert torch.equal(s2, send_ids) and torch.equal(r2, src) and torch.equal(i2, inv) and int(status[0]) == lost
        for o in range(world):
            assert torch.equal(lists[o][rank * C:(rank + 1) * C], send_ids[o * C:(o + 1) * C])
            rest = torch.cat([lists[o][:rank * C], lists[o][(rank + 1) * C:]])
            assert (rest == 3).all()


@pytest.mark.gpu
def test_slot_grads_push_matches_the_sample_order_delivery():
    """dctr_slot_grads_push (bucket order) writes exactly what the per-slot gradient of a model without a cross layer is:
    wts * dz[n, f*d .. f*d+d) at row rank * C + k of the owner's buffer; padded columns zero, unused rows untouched."""
    import deepctr_b200.ops  # noqa: F401
    O = torch.ops.deepctr
    from deepctr_b200 import _lib as L
    g = torch.Generator(device="cuda").manual_seed(5)
    for world, rank, N, F, d, weighted in ((2, 1, 300, 7, 11, True), (4, 0, 1025, 39, 10, False), (3, 2, 64, 5, 4, True)):
        M, ldg = N * F, (d + 3) // 4 * 4
        C = int(M / world * 1.3) + 8
        ids = torch.randint(0, 5000, (M,), device="cuda", dtype=torch.int32, generator=g)
        owner = torch.remainder(ids, world).to(torch.int32)
        owner_sorted, perm = torch.empty_like(ids), torch.empty_like(ids)
        ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
        O.sort_pairs(owner, None, M, world, owner_sorted, perm, ws)
        send_ids = torch.empty(world * C, dtype=torch.int32, device="cuda")
        src = torch.empty(world * C, dtype=torch.int32, device="cuda")
        inv = torch.empty(M, dtype=torch.int32, device="cuda")
        status = torch.zeros(2, dtype=torch.int32, device="cuda")
        O.shard_pack(ids, perm, owner_sorted, world, C, send_ids, src, inv, status)
        assert int(status[0]) == 0
        dz = torch.randn(N, F * d + 3, device="cuda", generator=g)[:, : F * d + 1]          # a strided view, bias column last
        wts = torch.rand(M, device="cuda", generator=g) if weighted else None
        bufs = [torch.full((world * C, ldg), 9.0, device="cuda") for _ in range(world)]
        peers = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device="cuda")
        O.slot_grads_push(send_ids, src, wts, dz, F, d, ldg, C, world, rank, peers, bufs[rank])
        torch.cuda.synchronize()
        rows = dz[:, : F * d].reshape(M, d) * (wts[:, None] if weighted else 1.0)
        for o in range(world):
            want = torch.full((world * C, ldg), 9.0, device="cuda")
            valid = torch.nonzero(send_ids[o * C:(o + 1) * C] >= 0).flatten()
            blk = torch.zeros(valid.numel(), ldg, device="cuda")
            blk[:, :d] = rows[src[o * C + valid].long()]
            want[rank * C + valid] = blk
            assert torch.equal(bufs[o], want), (world, rank, o)

