"""GPU parity tests, kernel by kernel, through the C-ABI (ctypes) against the numpy oracle.
Integer / index work (sort order, segment structure) is checked BIT-EXACT; fp32 kernels against the fp64
oracle at rtol 1e-5 (+ a small absolute floor), the tolerance BASELINE.json's north_star states."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from deepctr_b200 import _lib
    _lib.lib()
    return _lib


_KEEP = []   # device tensors stay referenced until the test ends: kernels are asynchronous and the caching
             # allocator would hand a dropped temporary's block to the next allocation


@pytest.fixture(autouse=True)
def _release():
    yield
    torch.cuda.synchronize()
    del _KEEP[:]


def dev(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    t = t.cuda()
    _KEEP.append(t)
    return t


def padded_table(table, ld):
    X, d = table.shape
    t = np.zeros((X, ld), np.float32)
    t[:, :d] = table
    return dev(t)


def r4(n):
    return (n + 3) // 4 * 4


def close(a, b, rtol=RTOL, atol=ATOL):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("cross,bias_first,act0,F,d,N", [
    ("none", True, "none", 16, 11, 100), ("none", False, "tanh", 39, 11, 33), ("ip", False, "relu", 16, 11, 65),
    ("ip", False, "tanh", 39, 11, 17), ("op", False, "sigmoid", 16, 11, 40), ("op", False, "tanh", 7, 3, 9),
    ("ip", False, "tanh", 5, 40, 12), ("none", True, "none", 1, 1, 3)])
def test_gather_concat_fwd(L, cross, bias_first, act0, F, d, N):
    import deepctr_oracle as O
    rng = np.random.default_rng(7)
    X, ld = 1000, r4(d)
    table = rng.normal(0, 0.3, (X, d)).astype(np.float32)
    ids = rng.integers(0, X, (N, F))
    ids[N // 2] = ids[0]
    wts = rng.uniform(0, 1.5, (N, F)).astype(np.float32)
    e = O.gather_fields(table.astype(np.float64), ids, wts.astype(np.float64))
    z = O.activate(act0, O.build_z(e, cross, 0.37, bias_first))
    D0 = z.shape[1]
    ldz = r4(D0) + 4
    out = torch.full((N, ldz), 7.0, device="cuda")
    bias = dev(np.array([0.37], np.float32))
    L.call("dctr_gather_concat_fwd", dev(ids, torch.int32).data_ptr(), dev(wts).data_ptr(), padded_table(table, ld).data_ptr(),
           X, ld, N, F, d, L.CROSS[cross], int(bias_first), bias.data_ptr(), L.ACT[act0], out.data_ptr(), ldz, None)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    close(got[:, :D0], z)
    assert (got[:, D0:] == 0).all()


def test_arrange_fields_bit_exact(L):
    """slot of a feature = its field; later feature of the same field overwrites; absent fields -> (pad_row, 0)
    (FNN_wnzh.py:91-95).  Index work: checked bit-exact against the host restatement."""
    rng = np.random.default_rng(31)
    N, J, F, X = 257, 16, 16, 5000
    feat_field = rng.integers(0, F, X).astype(np.int32)
    feats = rng.integers(0, X, (N, J)).astype(np.int32)
    feats[rng.random((N, J)) < 0.2] = -1
    want_ids = np.full((N, F), X, np.int32)
    want_wts = np.zeros((N, F), np.float32)
    for n in range(N):
        for j in range(J):
            if feats[n, j] >= 0:
                want_ids[n, feat_field[feats[n, j]]] = feats[n, j]
                want_wts[n, feat_field[feats[n, j]]] = 1.0
    ids = torch.zeros(N, F, dtype=torch.int32, device="cuda")
    wts = torch.full((N, F), 9.0, device="cuda")
    L.call("dctr_arrange_fields", dev(feats).data_ptr(), dev(feat_field).data_ptr(), N, J, F, X, ids.data_ptr(), wts.data_ptr(), None)
    torch.cuda.synchronize()
    assert np.array_equal(ids.cpu().numpy(), want_ids)
    assert np.array_equal(wts.cpu().numpy(), want_wts)


@pytest.mark.parametrize("M,X,ld", [(1, 3, 4), (1000, 37, 12), (70001, 5000, 12), (513, 9, 52), (64, 64, 8)])
def test_gather_rows_bit_exact(L, M, X, ld):
    """the owner-side lookup of the row-sharded table and the gradient permutation: a pure row copy (bit-exact)"""
    rng = np.random.default_rng(M)
    table = rng.normal(0, 1, (X, ld)).astype(np.float32)
    ids = rng.integers(0, X, M).astype(np.int32)
    out = torch.empty(M, ld, dtype=torch.float32, device="cuda")
    L.call("dctr_gather_rows", L.ptr(dev(ids)), L.ptr(dev(table)), X, ld, M, L.ptr(out), 0)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), table[ids])


@pytest.mark.parametrize("cross,F,d,N", [("none", 16, 11, 50), ("ip", 16, 11, 41), ("ip", 39, 11, 9), ("op", 16, 11, 23), ("op", 6, 5, 8)])
def test_cross_bwd(L, cross, F, d, N):
    import deepctr_oracle as O
    rng = np.random.default_rng(11)
    X, ld = 500, r4(d)
    table = rng.normal(0, 0.3, (X, d)).astype(np.float32)
    ids = rng.integers(0, X, (N, F))
    wts = rng.uniform(0.2, 1.5, (N, F)).astype(np.float32)
    cw = {"none": 0, "ip": F * (F - 1) // 2, "op": d * d}[cross]
    D0 = F * d + cw + 1
    ldz = r4(D0)
    dz = rng.normal(0, 1, (N, ldz)).astype(np.float32)
    e = O.gather_fields(table.astype(np.float64), ids, wts.astype(np.float64))
    de = dz[:, :F * d].astype(np.float64).reshape(N, F, d) + O.cross_backward(e, cross, dz[:, F * d:F * d + cw].astype(np.float64))
    want = de * wts[..., None]
    ge = torch.full((N, F, ld), 5.0, device="cuda")
    L.call("dctr_cross_bwd", dev(ids, torch.int32).data_ptr(), dev(wts).data_ptr(), padded_table(table, ld).data_ptr(), X, ld,
           N, F, d, L.CROSS[cross], dev(dz).data_ptr(), ldz, ge.data_ptr(), ld, None)
    torch.cuda.synchronize()
    got = ge.cpu().numpy()
    close(got[..., :d], want, atol=1e-5)
    assert (got[..., d:] == 0).all()


@pytest.mark.parametrize("F,d,N", [(16, 11, 100), (39, 3, 7), (16, 1, 33), (4, 51, 19)])
def test_fm_fwd(L, F, d, N):
    import deepctr_oracle as O
    rng = np.random.default_rng(3)
    X, ld = 800, r4(d)
    table = rng.normal(0, 0.2, (X, d)).astype(np.float32)
    ids = rng.integers(0, X, (N, F))
    ids[:, -1] = X - 1
    wts = rng.uniform(0, 1.5, (N, F)).astype(np.float32)
    wts[:, -1] = 0            # padded slot (ipinyou.py:58-59)
    y = rng.integers(0, 2, N).astype(np.float32)
    fm = O.OracleFM(table, 0.05)
    z, vx = fm.logits(ids, wts)
    f = lambda: torch.zeros(N, device="cuda")
    logit, pred, loss_n, g = f(), f(), f(), f()
    s = torch.full((N, ld), 3.0, device="cuda")
    L.call("dctr_fm_fwd", dev(ids, torch.int32).data_ptr(), dev(wts).data_ptr(), padded_table(table, ld).data_ptr(), X, ld, N, F, d,
           dev(np.array([0.05], np.float32)).data_ptr(), dev(y).data_ptr(), 0.5, logit.data_ptr(), pred.data_ptr(),
           loss_n.data_ptr(), g.data_ptr(), s.data_ptr(), ld, None)
    torch.cuda.synchronize()
    close(logit.cpu(), z)
    close(pred.cpu(), O.sigmoid(z))
    close(loss_n.cpu(), O.xent_with_logits(z, y))
    close(g.cpu(), (O.sigmoid(z) - y) * 0.5)
    close(s.cpu().numpy()[:, 1:d], vx)
    assert (s.cpu().numpy()[:, d:] == 0).all()


@pytest.mark.parametrize("M,X", [(1, 5), (31, 3), (2048, 1000), (2049, 70000), (65536, 1000001), (300000, 2 ** 27), (5000, 1)])
def test_sort_rows_bit_exact(L, M, X):
    rng = np.random.default_rng(M)
    ids = np.minimum((rng.pareto(1.1, M) * 3).astype(np.int64), X - 1) if X > 1 else np.zeros(M, np.int64)
    ids = (ids * 2654435761) % X
    order = np.argsort(ids, kind="stable")
    rows = torch.zeros(M, dtype=torch.int32, device="cuda")
    pos = torch.zeros(M, dtype=torch.int32, device="cuda")
    ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    L.call("dctr_sort_rows", dev(ids, torch.int32).data_ptr(), M, X, rows.data_ptr(), pos.data_ptr(), ws.data_ptr(), ws.numel(), None)
    torch.cuda.synchronize()
    assert np.array_equal(rows.cpu().numpy(), ids[order])
    assert np.array_equal(pos.cpu().numpy(), order)        # stability: equal rows keep (sample, field) order


def _sorted(L, ids, X):
    M = ids.size
    rows = torch.zeros(M, dtype=torch.int32, device="cuda")
    pos = torch.zeros(M, dtype=torch.int32, device="cuda")
    ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    L.call("dctr_sort_rows", dev(ids.reshape(-1), torch.int32).data_ptr(), M, X, rows.data_ptr(), pos.data_ptr(), ws.data_ptr(), ws.numel(), None)
    return rows, pos


@pytest.mark.parametrize("mode", ["sum", "seq"])
@pytest.mark.parametrize("N,F,d,X,hot", [(100, 16, 11, 300, False), (700, 16, 11, 50, True), (64, 39, 11, 2000, True), (33, 3, 20, 10, False)])
def test_scatter_sgd(L, mode, N, F, d, X, hot):
    import deepctr_oracle as O
    rng = np.random.default_rng(5)
    ld = r4(d)
    table = rng.normal(0, 0.5, (X, d)).astype(np.float32)
    ids = rng.integers(0, X, (N, F))
    if hot:
        ids[:, 0] = 1          # one row shared by every sample (Criteo numeric rows): long-segment path
        ids[: N // 2, 1] = 2
    wts = rng.uniform(0.2, 1.5, (N, F)).astype(np.float32)
    ldz = r4(1 + F * d)
    dz = rng.normal(0, 1, (N, ldz)).astype(np.float32)
    ge = dz[:, 1:1 + F * d].astype(np.float64).reshape(N, F, d) * wts[..., None]
    lr, decay = 0.01, 0.999
    want = table.astype(np.float64)
    if mode == "seq":
        O.seq_decay_scatter(want, ids, ge, lr, decay)
    else:
        upd = np.zeros_like(want)
        np.add.at(upd, ids.reshape(-1), ge.reshape(-1, d))
        touched = np.unique(ids)
        want[touched] = want[touched] * decay - lr * upd[touched]
    rows, pos = _sorted(L, ids, X)
    t = padded_table(table, ld)
    ws = torch.zeros(L.lib().dctr_scatter_workspace(N * F, ld), dtype=torch.uint8, device="cuda")
    dzd = dev(dz)
    L.call("dctr_scatter_sgd", L.SCATTER_SEQ_DECAY if mode == "seq" else L.SCATTER_SUM, rows.data_ptr(), pos.data_ptr(), N * F, F, d,
           dzd.data_ptr() + 4, ldz, d, dev(wts).data_ptr(), lr, decay, t.data_ptr(), X, ld, ws.data_ptr(), ws.numel(), None)
    torch.cuda.synchronize()
    got = t.cpu().numpy()
    close(got[:, :d], want, atol=2e-6)
    assert (got[:, d:] == 0).all()
    untouched = np.setdiff1d(np.arange(X), ids)
    assert np.array_equal(got[untouched, :d], table[untouched])     # touched-row set is exact


def test_scatter_deterministic(L):
    rng = np.random.default_rng(9)
    N, F, d, X, ld = 2000, 16, 11, 100, 12
    ids = rng.integers(0, X, (N, F))
    dz = dev(rng.normal(0, 1, (N, F * d + 3)).astype(np.float32))
    table = rng.normal(0, 0.5, (X, ld)).astype(np.float32)
    outs = []
    for _ in range(2):
        rows, pos = _sorted(L, ids, X)
        t = dev(table)
        ws = torch.zeros(L.lib().dctr_scatter_workspace(N * F, ld), dtype=torch.uint8, device="cuda")
        L.call("dctr_scatter_sgd", L.SCATTER_SUM, rows.data_ptr(), pos.data_ptr(), N * F, F, d, dz.data_ptr(), F * d + 3, d, None,
               0.1, 1.0, t.data_ptr(), X, ld, ws.data_ptr(), ws.numel(), None)
        torch.cuda.synchronize()
        outs.append(t.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])


@pytest.mark.parametrize("lam", [0.0, 0.01])
def test_fm_scatter_and_sweep(L, lam):
    import deepctr_oracle as O
    rng = np.random.default_rng(13)
    N, F, d, X = 300, 16, 11, 400
    ld = r4(d)
    table = rng.normal(0, 0.2, (X, d)).astype(np.float32)
    ids = rng.integers(0, X - 100, (N, F))
    ids[:, 0] = 7
    wts = rng.uniform(0, 1.5, (N, F)).astype(np.float32)
    y = rng.integers(0, 2, N).astype(np.float32)
    lr = 0.05
    fm = O.OracleFM(table, 0.0)
    fm.train_step(ids, wts, y, lr, lam)
    t = padded_table(table, ld)
    idsd, wtsd = dev(ids, torch.int32), dev(wts)
    g, s = torch.zeros(N, device="cuda"), torch.zeros(N, ld, device="cuda")
    L.call("dctr_fm_fwd", idsd.data_ptr(), wtsd.data_ptr(), t.data_ptr(), X, ld, N, F, d, None, dev(y).data_ptr(), 1.0,
           None, None, None, g.data_ptr(), s.data_ptr(), ld, None)
    rows, pos = _sorted(L, ids, X)
    ws = torch.zeros(L.lib().dctr_scatter_workspace(N * F, ld), dtype=torch.uint8, device="cuda")
    decay = 1 - lr * lam
    L.call("dctr_scatter_sgd_fm", rows.data_ptr(), pos.data_ptr(), N * F, F, d, g.data_ptr(), s.data_ptr(), ld, wtsd.data_ptr(),
           lr, decay, t.data_ptr(), X, ld, ws.data_ptr(), ws.numel(), None)
    if lam:
        L.call("dctr_decay_sweep", t.data_ptr(), X, ld, decay, rows.data_ptr(), N * F, None)
    torch.cuda.synchronize()
    close(t.cpu().numpy()[:, :d], fm.table, atol=2e-6)


GEMMS = ["dctr_gemm", "dctr_gemm_tc"]     # fp32 FFMA path and tcgen05 3xTF32 path: same contract, same tolerance


@pytest.mark.parametrize("fn", GEMMS)
@pytest.mark.parametrize("layout", ["nn", "nt", "tn"])
@pytest.mark.parametrize("M,N,K", [(100, 300, 177), (4096, 100, 300), (33, 7, 5), (257, 129, 1171), (64, 64, 16), (1, 1, 1),
                                   (300, 260, 70), (33000, 135, 40)])      # the last one: > 16 MB of output = the transposing epilogue
def test_gemm_layouts(L, fn, layout, M, N, K):
    rng = np.random.default_rng(17)
    A = rng.normal(0, 1, (M, K)).astype(np.float32)
    B = rng.normal(0, 1, (K, N)).astype(np.float32)
    want = A.astype(np.float64) @ B.astype(np.float64)

    def pad(a):
        r, c = a.shape
        o = np.zeros((r, r4(c) + 4), np.float32)
        o[:, :c] = a
        return dev(o)

    Ad = pad(A if layout != "tn" else A.T.copy())
    Bd = pad(B if layout != "nt" else B.T.copy())
    ldc = r4(N) + 4
    splits = 3 if layout == "tn" else 1
    C = torch.full((splits, M, ldc), 9.0, device="cuda")
    lay = {"nn": L.GEMM_NN, "nt": L.GEMM_NT, "tn": L.GEMM_TN}[layout]
    L.call(fn, lay, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0), C.data_ptr(), ldc,
           L.EPI_NONE, 0, None, None, 0, None, 0, splits, None)
    if splits > 1:
        out = torch.zeros(M, ldc, device="cuda")
        L.call("dctr_splitk_reduce", C.data_ptr(), splits, M, N, ldc, out.data_ptr(), ldc, None)
    else:
        out = C[0]
    torch.cuda.synchronize()
    got = out.cpu().numpy()[:, :N]
    scale = np.sqrt(K)
    # tcgen05 accumulates with truncation: its stated absolute floor is 1.5x the FFMA path's
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=(3e-6 if fn == "dctr_gemm_tc" else 2e-6) * scale)


@pytest.mark.parametrize("fn", GEMMS)
@pytest.mark.parametrize("act", ["tanh", "relu", "sigmoid", "none"])
@pytest.mark.parametrize("bcast", [False, True])
@pytest.mark.parametrize("M,N,K", [(130, 70, 45), (33000, 135, 45)])       # small: row-per-lane stores; > 16 MB: transposing epilogue
def test_gemm_epilogues(L, fn, act, bcast, M, N, K):
    import deepctr_oracle as O
    rng = np.random.default_rng(19)
    A = rng.normal(0, 0.5, (M, r4(K))).astype(np.float32); A[:, K:] = 0
    B = rng.normal(0, 0.5, (K, r4(N))).astype(np.float32); B[:, N:] = 0
    bias = rng.normal(0, 0.5, N).astype(np.float32)
    mrows = 1 if bcast else M
    mask = ((rng.random((mrows, r4(N))) < 0.6) / (1.0 if bcast else 0.6)).astype(np.float32)
    pre = A[:, :K].astype(np.float64) @ B[:, :N].astype(np.float64) + bias
    want = O.activate(act, pre) * mask[:, :N]
    C = torch.zeros(M, r4(N), device="cuda")
    md = dev(mask)
    L.call(fn, L.GEMM_NN, M, N, K, dev(A).data_ptr(), r4(K), dev(B).data_ptr(), r4(N), C.data_ptr(), r4(N),
           L.EPI_FWD, L.ACT[act], dev(bias).data_ptr(), None, 0, md.data_ptr(), 0 if bcast else r4(N), 1, None)
    torch.cuda.synchronize()
    close(C.cpu().numpy()[:, :N], want, atol=3e-6)
    # DGRAD: dA = (dY @ W^T) * mask * act'(stored/mask); stored = the forward output `want`-like tensor
    dY = rng.normal(0, 1, (M, r4(K))).astype(np.float32); dY[:, K:] = 0          # [M, K] plays dY, W = B^T-shaped [N, K]
    W = rng.normal(0, 0.5, (N, r4(K))).astype(np.float32); W[:, K:] = 0
    stored = np.zeros((M, r4(N)), np.float32); stored[:, :N] = want
    din = dY[:, :K].astype(np.float64) @ W[:, :K].astype(np.float64).T
    mk = np.broadcast_to(mask[:, :N], (M, N)).astype(np.float64)
    t = np.where(mk != 0, stored[:, :N] / np.where(mk != 0, mk, 1), 0)
    want_d = din * mk * O.activate_grad_from_out(act, t)
    D = torch.zeros(M, r4(N), device="cuda")
    L.call(fn, L.GEMM_NT, M, N, K, dev(dY).data_ptr(), r4(K), dev(W).data_ptr(), r4(K), D.data_ptr(), r4(N),
           L.EPI_DGRAD, L.ACT[act], None, dev(stored).data_ptr(), r4(N), md.data_ptr(), 0 if bcast else r4(N), 1, None)
    torch.cuda.synchronize()
    close(D.cpu().numpy()[:, :N], want_d, atol=1e-5)


def test_output_unit_and_reductions(L):
    import deepctr_oracle as O
    rng = np.random.default_rng(23)
    N, H = 1000, 100
    a = rng.normal(0, 1, (N, r4(H))).astype(np.float32)
    w = rng.normal(0, 0.3, H).astype(np.float32)
    y = rng.integers(0, 2, N).astype(np.float32)
    z = a[:, :H].astype(np.float64) @ w + 0.2
    f = lambda: torch.zeros(N, device="cuda")
    logit, pred, loss_n, dl = f(), f(), f(), f()
    ad, wd = dev(a), dev(w)
    ws = torch.zeros(L.lib().dctr_reduce_workspace(N, H), dtype=torch.uint8, device="cuda")
    lsum = torch.zeros(1, device="cuda")
    for rep in range(2):        # twice: the ticket workspace must come back ready for the next launch
        L.call("dctr_out_xent_fwd", ad.data_ptr(), r4(H), N, H, wd.data_ptr(), dev(np.array([0.2], np.float32)).data_ptr(),
               dev(y).data_ptr(), 1.0, logit.data_ptr(), pred.data_ptr(), loss_n.data_ptr(), dl.data_ptr(), 0.5, lsum.data_ptr(),
               ws.data_ptr(), ws.numel(), None)
        close(logit.cpu(), z, atol=1e-5)
        close(loss_n.cpu(), O.xent_with_logits(z, y), atol=1e-5)
        close(dl.cpu(), O.sigmoid(z) - y, atol=1e-5)
        close(lsum.cpu(), [0.5 * O.xent_with_logits(z, y).sum()], atol=1e-4)
    out = torch.zeros(1, device="cuda")
    L.call("dctr_vec_sum", loss_n.data_ptr(), N, 0.5, out.data_ptr(), None)
    close(out.cpu(), [0.5 * loss_n.cpu().numpy().astype(np.float64).sum()])
    gw = torch.zeros(H, device="cuda")
    want_gw = (dl.cpu().numpy().astype(np.float64)[:, None] * a[:, :H]).sum(0)
    for rep in range(2):
        L.call("dctr_col_wsum", ad.data_ptr(), r4(H), N, H, dl.data_ptr(), gw.data_ptr(), ws.data_ptr(), ws.numel(), None)
        close(gw.cpu(), want_gw, atol=1e-4)
    da = torch.zeros(N, r4(H), device="cuda")
    L.call("dctr_out_bwd", dl.data_ptr(), wd.data_ptr(), ad.data_ptr(), r4(H), N, H, L.ACT["tanh"], None, 0, da.data_ptr(), r4(H), None)
    want_da = dl.cpu().numpy().astype(np.float64)[:, None] * w * (1 - a[:, :H].astype(np.float64) ** 2)
    close(da.cpu().numpy()[:, :H], want_da, atol=1e-5)
    # fused: da + gw + gb in one launch, with a dropout-style mask
    mask = ((rng.random((N, r4(H))) < 0.7) / 0.7).astype(np.float32)
    am = (a * mask).astype(np.float32)                      # stored forward output = act * mask
    amd, md = dev(am), dev(mask)
    da2, gw2, gb2 = torch.zeros(N, r4(H), device="cuda"), torch.zeros(H, device="cuda"), torch.zeros(1, device="cuda")
    for rep in range(2):
        L.call("dctr_out_bwd_fused", dl.data_ptr(), wd.data_ptr(), amd.data_ptr(), r4(H), N, H, L.ACT["tanh"], md.data_ptr(), r4(H),
               da2.data_ptr(), r4(H), gw2.data_ptr(), gb2.data_ptr(), ws.data_ptr(), ws.numel(), None)
        dl64 = dl.cpu().numpy().astype(np.float64)
        close(gw2.cpu(), (dl64[:, None] * am[:, :H]).sum(0), atol=1e-4)
        close(gb2.cpu(), [dl64.sum()], atol=1e-4)
        close(da2.cpu().numpy()[:, :H], dl64[:, None] * w * mask[:, :H] * (1 - a[:, :H].astype(np.float64) ** 2) * (mask[:, :H] != 0), atol=1e-5)
    p = dev(rng.normal(0, 1, 1001).astype(np.float32)); g = dev(rng.normal(0, 1, 1001).astype(np.float32))
    want = p.cpu().numpy().astype(np.float64) * 0.99 - 0.1 * g.cpu().numpy()
    L.call("dctr_dense_sgd", p.data_ptr(), g.data_ptr(), 1001, 0.1, 0.99, None)
    torch.cuda.synchronize()
    close(p.cpu(), want)


@pytest.mark.parametrize("N,H,off,pad", [(1, 1, 0, 0), (37, 50, 0, 2), (4096, 300, 0, 0), (5000, 1, 177, 3), (70000, 400, 0, 0),
                                          (3000, 129, 1, 4), (9000, 1000, 0, 0), (260, 7, 0, 0)])
def test_column_reductions_shapes(L, N, H, off, pad):
    """dctr_col_wsum / dctr_out_bwd_fused over the shapes the engine uses: wide and narrow, 16-byte aligned and not (a
    single column inside a wider matrix), N far above and below one row chunk; run twice (ticket reset) and identical
    both times (fixed-order reduction)."""
    rng = np.random.default_rng(N + H)
    lda = r4(H + off) + 4 * (pad // 2) if off == 0 else H + off + pad
    A = rng.normal(0, 1, (N, lda)).astype(np.float32)
    wv = rng.normal(0, 1, N).astype(np.float32)
    Ad, wd = dev(A), dev(wv)
    ws = torch.zeros(L.lib().dctr_reduce_workspace(N, H), dtype=torch.uint8, device="cuda")
    out = torch.zeros(H, device="cuda")
    want = (wv.astype(np.float64)[:, None] * A[:, off:off + H]).sum(0)
    tol = 2e-6 * np.sqrt(N) * 4
    res = []
    for rep in range(2):
        L.call("dctr_col_wsum", Ad.data_ptr() + 4 * off, lda, N, H, wd.data_ptr(), out.data_ptr(), ws.data_ptr(), ws.numel(), None)
        res.append(out.cpu().numpy().copy())
        close(res[-1], want, atol=tol)
    assert np.array_equal(res[0], res[1])
    L.call("dctr_col_wsum", Ad.data_ptr() + 4 * off, lda, N, H, None, out.data_ptr(), ws.data_ptr(), ws.numel(), None)
    close(out.cpu(), A[:, off:off + H].astype(np.float64).sum(0), atol=tol)
    # fused output-unit backward; row-broadcast mask (ldmask = 0, the FNN_wnzh dropout form) and none
    w = rng.normal(0, 0.3, H).astype(np.float32)
    act_out = np.tanh(A[:, off:off + H].astype(np.float64))
    for use_mask in (False, True):
        mask = ((rng.random(r4(H)) < 0.6) * 2.0).astype(np.float32) if use_mask else None
        stored = np.zeros((N, lda), np.float32)
        stored[:, off:off + H] = act_out * (mask[:H] if use_mask else 1.0)
        sd = dev(stored)
        da = torch.full((N, lda), 7.0, device="cuda")
        gw, gb = torch.zeros(H, device="cuda"), torch.zeros(1, device="cuda")
        L.call("dctr_out_bwd_fused", wd.data_ptr(), dev(w).data_ptr(), sd.data_ptr() + 4 * off, lda, N, H, L.ACT["tanh"],
               dev(mask).data_ptr() if use_mask else None, 0, da.data_ptr() + 4 * off, lda, gw.data_ptr(), gb.data_ptr(),
               ws.data_ptr(), ws.numel(), None)
        wv64 = wv.astype(np.float64)
        close(gw.cpu(), (wv64[:, None] * stored[:, off:off + H]).sum(0), atol=tol)
        close(gb.cpu(), [wv64.sum()], atol=tol)
        m = mask[:H].astype(np.float64) if use_mask else np.ones(H)
        dan = da.cpu().numpy()
        close(dan[:, off:off + H], wv64[:, None] * w * m * (1 - act_out ** 2), atol=1e-5)
        assert np.all(dan[:, :off] == 7.0) and np.all(dan[:, off + H:] == 7.0)       # nothing written outside the H columns


@pytest.mark.parametrize("N,ties,pos_rate", [(1, False, 1.0), (2, False, 0.5), (1000, False, 0.2), (5000, True, 0.05),
                                              (70001, True, 0.5), (300, False, 0.0), (4096, True, 1.0)])
def test_eval_metrics_vs_sklearn(L, N, ties, pos_rate):
    """AUC (ties by average rank), log-loss and RMSE on the device equal sklearn's on the host (baseline.py:246-259)"""
    from sklearn.metrics import log_loss, mean_squared_error, roc_auc_score
    from deepctr_b200.metrics import device_metrics
    rng = np.random.default_rng(N)
    y = (rng.random(N) < pos_rate).astype(np.float32)
    if 0.0 < pos_rate < 1.0 and N > 1:
        y[0], y[1] = 1.0, 0.0
    p = (1.0 / (1.0 + np.exp(-(rng.normal(0, 1.5, N) + 1.5 * y)))).astype(np.float32)
    if ties:
        p = np.round(p, 2).astype(np.float32)               # heavy ties, some exactly 0 or 1
    got = [device_metrics(p, y) for _ in range(2)]
    assert got[0] == got[1]                                    # deterministic
    g = got[0]
    assert g["n_pos"] == int(y.sum())
    if y.min() == y.max():
        assert g["auc"] == -1.0
    else:
        np.testing.assert_allclose(g["auc"], roc_auc_score(y, p), rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(g["log_loss"], log_loss(y, np.clip(p.astype(np.float64), 1e-15, 1 - 1e-15), labels=[0, 1]), rtol=1e-9)
    np.testing.assert_allclose(g["rmse"], np.sqrt(mean_squared_error(y, p.astype(np.float64))), rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("N,H,mask_kind", [(1, 1, None), (37, 50, "row"), (4096, 100, "row"), (5000, 129, "full"), (70000, 400, None),
                                           (300, 1000, "full"), (2049, 1024, None)])
def test_out_unit_train_equals_two_kernel_path(L, N, H, mask_kind):
    """dctr_out_unit_train == dctr_out_xent_fwd + dctr_out_bwd_fused (each checked against the oracle above): same logit /
    pred / loss / dlogit bit for bit (same GEMV order is not promised: rtol 1e-6), da, gw, gb at reduction tolerance"""
    rng = np.random.default_rng(N * 7 + H)
    ld = r4(H)
    act_out = np.tanh(rng.normal(0, 1, (N, ld))).astype(np.float32)
    mask = None
    if mask_kind == "row":
        mask = ((rng.random((1, ld)) < 0.6) * 2.0).astype(np.float32)
    elif mask_kind == "full":
        mask = ((rng.random((N, ld)) < 0.7) / 0.7).astype(np.float32)
    stored = act_out * (mask if mask is not None else 1.0)
    stored[:, H:] = 0
    w = rng.normal(0, 0.3, H).astype(np.float32)
    y = rng.integers(0, 2, N).astype(np.float32)
    sd, wd, yd, bd = dev(stored.astype(np.float32)), dev(w), dev(y), dev(np.array([0.1], np.float32))
    md = dev(mask) if mask is not None else None
    ldm = 0 if mask_kind == "row" else ld
    ws = torch.zeros(L.lib().dctr_reduce_workspace(N, H + 1), dtype=torch.uint8, device="cuda")
    z = lambda *s: torch.zeros(*s, device="cuda")
    ref = dict(logit=z(N), pred=z(N), loss_n=z(N), dl=z(N), loss=z(1), da=z(N, ld), gw=z(H), gb=z(1))
    L.call("dctr_out_xent_fwd", sd.data_ptr(), ld, N, H, wd.data_ptr(), bd.data_ptr(), yd.data_ptr(), 0.5, ref["logit"].data_ptr(),
           ref["pred"].data_ptr(), ref["loss_n"].data_ptr(), ref["dl"].data_ptr(), 0.5, ref["loss"].data_ptr(), ws.data_ptr(), ws.numel(), None)
    L.call("dctr_out_bwd_fused", ref["dl"].data_ptr(), wd.data_ptr(), sd.data_ptr(), ld, N, H, L.ACT["tanh"], L.ptr(md), ldm,
           ref["da"].data_ptr(), ld, ref["gw"].data_ptr(), ref["gb"].data_ptr(), ws.data_ptr(), ws.numel(), None)
    runs = []
    for rep in range(2):
        got = dict(logit=z(N), pred=z(N), loss_n=z(N), dl=z(N), loss=z(1), da=z(N, ld), gw=z(H), gb=z(1))
        L.call("dctr_out_unit_train", sd.data_ptr(), ld, N, H, wd.data_ptr(), bd.data_ptr(), yd.data_ptr(), 0.5, L.ACT["tanh"],
               L.ptr(md), ldm, got["logit"].data_ptr(), got["pred"].data_ptr(), got["loss_n"].data_ptr(), got["dl"].data_ptr(), 0.5,
               got["loss"].data_ptr(), got["da"].data_ptr(), ld, got["gw"].data_ptr(), got["gb"].data_ptr(), ws.data_ptr(), ws.numel(), None)
        torch.cuda.synchronize()
        runs.append({k: v.cpu().numpy().copy() for k, v in got.items()})
    for k in runs[0]:
        assert np.array_equal(runs[0][k], runs[1][k]), k            # deterministic, ticket reset
    tol = 2e-6 * np.sqrt(N) * 4
    for k in ("logit", "pred", "loss_n", "dl"):
        close(runs[0][k], ref[k].cpu(), rtol=2e-6, atol=2e-6)
    close(runs[0]["loss"], ref["loss"].cpu(), rtol=1e-5, atol=tol)
    close(runs[0]["da"][:, :H], ref["da"].cpu().numpy()[:, :H], rtol=1e-5, atol=1e-6)
    close(runs[0]["gw"], ref["gw"].cpu(), rtol=1e-5, atol=tol)
    close(runs[0]["gb"], ref["gb"].cpu(), rtol=1e-5, atol=tol)


def test_empty_inputs(L):
    # N = 0 / M = 0 are accepted and launch nothing
    before = L.launch_count()
    L.call("dctr_gather_concat_fwd", 16, None, 16, 10, 12, 0, 16, 11, 0, 1, None, 0, 16, 180, None)
    L.call("dctr_sort_rows", None, 0, 10, None, None, None, 0, None)
    L.call("dctr_gemm", L.GEMM_NN, 0, 4, 4, 16, 4, 16, 4, 16, 4, 0, 0, None, None, 0, None, 0, 1, None)
    assert L.launch_count() == before
