"""GPU parity of the SURVEY 8(f) widening rows through the C-ABI / the model classes against the oracle:
f1 FTRL (tf_util.py:22-24), f2 device-side text ingestion (FNN_wnzh.py:240-253, ipinyou.py:42-65, baseline.py:190-198),
f4 the SNN first layer as an embedding bag (SNN_DAE.py:234-260, 282-288)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from deepctr_b200 import _lib
    _lib.lib()
    return _lib


def close(a, b, rtol=1e-5, atol=1e-6):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


# ---- f1: FTRL ----------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,lr_power,l1,l2,gl2,zero", [(4, -0.5, 0.0, 0.0, 0.0, 0), (1000, -0.5, 0.0, 0.0, 0.01, 1),
                                                      (70000, -0.5, 0.05, 0.1, 0.0, 1), (260, -0.3, 0.01, 0.0, 0.0, 0)])
def test_dense_ftrl_kernel(L, n, lr_power, l1, l2, gl2, zero):
    """dctr_dense_ftrl against ftrl_apply over three steps; a third of the elements get no gradient (dense semantics:
    they are re-derived from `linear`, i.e. go to 0 on step one), tiny gradients exercise the cancellation-free sqrt form"""
    import deepctr_oracle as O
    rng = np.random.default_rng(n)
    lr = 0.05
    p = rng.uniform(-0.5, 0.5, n)
    slots = [np.full(n, 0.1), np.zeros(n)]
    tp = torch.tensor(p, dtype=torch.float32, device="cuda")
    tn = torch.full((n,), 0.1, dtype=torch.float32, device="cuda")
    tz = torch.zeros(n, dtype=torch.float32, device="cuda")
    for step in range(3):
        g = rng.normal(0, 1, n) * rng.choice([1.0, 1e-4], n)
        g[rng.random(n) < 0.33] = 0.0
        g = g.astype(np.float32).astype(np.float64)
        tg = torch.tensor(g, dtype=torch.float32, device="cuda")
        L.call("dctr_dense_ftrl", tp.data_ptr(), tg.data_ptr(), tn.data_ptr(), tz.data_ptr(), n, lr, lr_power, l1, l2, gl2, zero, None)
        p = O.ftrl_apply(p, g + gl2 * p, slots, lr, lr_power, l1, l2)
        close(tp.cpu(), p, 2e-5, 1e-6)
        close(tn.cpu(), slots[0], 2e-5, 1e-7)
        close(tz.cpu(), slots[1], 2e-5, 2e-6)
        assert bool((tg == 0).all()) == bool(zero) or not g.any()
    if l1 > 0:
        assert (p == 0).any() and (tp.cpu().numpy()[p == 0] == 0).all()      # the l1 threshold cuts the same elements


def test_dense_ftrl_rejects_bad_arguments(L):
    t = torch.zeros(8, dtype=torch.float32, device="cuda")
    a = t.data_ptr()
    for args in [(a, a, a, a, 6, 0.1, -0.5, 0.0, 0.0, 0.0, 0, None),           # n % 4
                 (a, a, a, a, 8, 0.0, -0.5, 0.0, 0.0, 0.0, 0, None),           # lr
                 (a, a + 4, a, a, 4, 0.1, -0.5, 0.0, 0.0, 0.0, 0, None),       # alignment
                 (a, None, a, a, 8, 0.1, -0.5, 0.0, 0.0, 0.0, 0, None)]:
        with pytest.raises(RuntimeError):
            L.call("dctr_dense_ftrl", *args)


@pytest.mark.parametrize("kind,lam,reduce,use_graph", [("fm", 0.0, "sum", False), ("fm", 1e-2, "mean", True), ("lr", 1e-2, "sum", True)])
def test_fm_lr_ftrl_steps(kind, lam, reduce, use_graph):
    """['ftrl', lr, reduce] for FM / LR (tf_util.py:22-24): dense FtrlOptimizer over W, V, b, L2 gradient inside g"""
    import deepctr_oracle as O
    from deepctr_b200.models import FM, LR
    rng = np.random.default_rng(515)
    N, X, F, rank = 64, 400, 8, 6
    init = ["uniform", -0.05, 0.05, [11, 12], None]
    if kind == "fm":
        m = FM(N, [X, F, rank], init, ["ftrl", 0.05, reduce], [lam], "train", 50)
    else:
        m = LR(N, [X, F], init, ["ftrl", 0.05, reduce], [lam], "train", 50)
    m.use_graph = use_graph
    o = O.OracleFM(m.engine.get_table(), 0.0)
    for step in range(6):
        ids = rng.integers(0, X // 2 if step >= 3 else X, (N, F))
        ids[:, 0] = 7
        wts = rng.uniform(0.5, 1.5, (N, F)).astype(np.float32)
        y = rng.integers(0, 2, N)
        loss, preds = m.train_step({m.sp_id_hldr: ids.reshape(-1), m.sp_wt_hldr: wts.reshape(-1), m.lbl_hldr: y})
        oloss, opreds = o.train_step_ftrl(ids, wts, y, 0.05, lam, reduce)
        close(preds, opreds, 5e-5)
        close(m.engine.get_table(), o.table, 1e-4, 2e-6)
        close(m.engine.bias().cpu(), [o.b], 1e-4, 2e-6)
    if lam == 0.0:
        close([loss], [oloss], 2e-5)


@pytest.mark.parametrize("gemm", ["simt", "tc"])
def test_fnn_ftrl_steps(tmp_path, monkeypatch, gemm):
    """['ftrl', lr, reduce] for the FNN family: table (densified gradient) + MLP under TF's FtrlOptimizer"""
    monkeypatch.setenv("DCTR_GEMM", gemm)
    from test_models_gpu import compare_state, fnn_batch, make_fnn
    rng = np.random.default_rng(616)
    N = 64
    model, net, (cat_sizes, offsets, n_num) = make_fnn("FNN", tmp_path, rng, N, [40, 30], "tanh", use_graph=(gemm == "tc"),
                                                       ptmzr=["ftrl", 0.05, "sum"])
    for step in range(5):
        v_wts, c_ids, c_wts, y, g_ids, g_wts = fnn_batch(rng, N, cat_sizes, offsets, n_num)
        loss, preds = model.train_step({model.v_wt_hldr: v_wts, model.c_id_hldr: c_ids, model.c_wt_hldr: c_wts, model.lbl_hldr: y})
        oloss, opreds = net.train_step_tf_ftrl(g_ids, g_wts, y, 0.05)
        close([loss], [oloss], 5e-5)
        close(preds, opreds, 1e-4)
        compare_state(model, net, 2e-4)


# ---- f2: device-side text ingestion ------------------------------------------------------------------------------------
def _fm_text(rng, n, F, X, ragged=True, crlf=False, blank_every=0):
    lines = []
    for i in range(n):
        k = int(rng.integers(1, F + 1)) if ragged else F
        ids = rng.integers(0, X, k)
        sep = "  " if i % 7 == 3 else " "
        lines.append("%d%s%s" % (rng.integers(0, 2), sep, sep.join("%d:%d" % (j, 1) for j in ids)))
        if blank_every and i % blank_every == 0:
            lines.append("")
    eol = "\r\n" if crlf else "\n"
    return eol.join(lines) + (eol if n % 2 else "")          # with and without a trailing newline


@pytest.mark.parametrize("n,F,X,crlf,blank_every", [(1, 16, 100, False, 0), (33, 16, 10 ** 6, False, 5), (1000, 16, 2 ** 31 - 1, True, 0),
                                                    (4097, 5, 50, False, 3)])
def test_index_and_parse_fm_lines_bit_exact(L, n, F, X, crlf, blank_every):
    """train.fm.txt lines (FNN_wnzh.py:248-252) parsed on the device == the reference parser, bit for bit: line count,
    ids in order, padding, labels; ragged lines, double spaces, CRLF, empty lines, missing final newline"""
    import text_oracle as T
    from deepctr_b200.ingest import DeviceText
    rng = np.random.default_rng(n)
    text = _fm_text(rng, n, F, X, crlf=crlf, blank_every=blank_every)
    lines = T.nonblank_lines(text)
    dt = DeviceText(text.encode())
    assert dt.n_lines == len(lines) == n
    starts = np.cumsum([0] + [len(s) + 1 for s in text.split("\n")])[:-1]
    starts = [int(s) for s, ln in zip(starts, text.split("\n")) if ln != ""]
    assert dt.line_start.cpu().numpy()[:n].tolist() == starts and int(dt.line_start[n]) == len(text)
    feats, wts, y, status = dt.parse_sparse(F, skip=0, pad_id=-1)
    of, ow, oy, onnz = T.sparse_block(lines, F, -1, 0)
    assert np.array_equal(feats.cpu().numpy(), of) and np.array_equal(wts.cpu().numpy(), ow)
    assert np.array_equal(y.cpu().numpy(), oy) and np.array_equal(status.cpu().numpy(), onnz)
    # fewer slots than features: truncated, status still reports the full count
    f2, _, _, s2 = dt.parse_sparse(2, skip=0, pad_id=7, with_wts=False)
    of2, _, _, _ = T.sparse_block(lines, 2, 7, 0)
    assert np.array_equal(f2.cpu().numpy(), of2) and np.array_equal(s2.cpu().numpy(), onnz)


def test_parse_yzx_and_long_lines(L):
    """*.yzx.txt (ipinyou.py:50-59): the token after y is skipped, rows are padded with (X_dim - 1, 0); lines longer than the
    24 KB shared-memory stage take the unstaged path; a capacity smaller than the file keeps the first max_lines lines"""
    import text_oracle as T
    from deepctr_b200.ingest import DeviceText
    rng = np.random.default_rng(5)
    lines = []
    for i in range(300):
        k = 3000 if i in (17, 18, 250) else int(rng.integers(0, 20))        # 3000 features ~ 30 KB
        lines.append("%d %d %s" % (rng.integers(0, 2), rng.integers(0, 300), " ".join("%d:1" % j for j in rng.integers(0, 10 ** 6, k))))
    text = "\n".join(lines) + "\n"
    dt = DeviceText(text.encode())
    J, pad = 24, 10 ** 6
    feats, wts, y, status = dt.parse_sparse(J, skip=1, pad_id=pad)
    of, ow, oy, onnz = T.sparse_block([ln.rstrip() for ln in lines], J, pad, 1)
    assert np.array_equal(feats.cpu().numpy(), of) and np.array_equal(wts.cpu().numpy(), ow)
    assert np.array_equal(y.cpu().numpy(), oy) and np.array_equal(status.cpu().numpy(), onnz)
    small = DeviceText(text.encode(), max_lines=100)
    assert small.n_lines == 300
    f3, _, y3, s3 = small.parse_sparse(J, skip=1, pad_id=pad)
    assert f3.shape[0] == 100 and np.array_equal(f3.cpu().numpy(), of[:100]) and np.array_equal(y3.cpu().numpy(), oy[:100])


def test_parse_flags_malformed_lines(L):
    from deepctr_b200.ingest import DeviceText, check_status
    dt = DeviceText(b"1 5:1 7:1\n   \n0 x:1\n1 99999999999:1\n0 3:1 4\n")
    feats, wts, y, status = dt.parse_sparse(4)
    assert status.cpu().numpy().tolist() == [2, -1, -2, -2, 2]
    assert feats.cpu().numpy()[0].tolist() == [5, 7, -1, -1] and feats.cpu().numpy()[4].tolist() == [3, 4, -1, -1]
    with pytest.raises(ValueError):
        check_status(status)
    empty = DeviceText(b"")
    assert empty.n_lines == 0 and empty.parse_sparse(4)[0].shape[0] == 0
    assert DeviceText(b"\n\n\n").n_lines == 0


@pytest.mark.parametrize("n", [1, 257, 5000])
def test_parse_criteo_tsv(L, n):
    """Criteo indexed TSV (clean.py:148-184) -> the engines' [N, 39] block == get_fxy (baseline.py:190-198): global ids bit-exact,
    numeric weights = float32(value / max_vals) (same double division), labels"""
    import text_oracle as T
    from deepctr_b200 import data as D
    from deepctr_b200.ingest import DeviceText, check_status
    rng = np.random.default_rng(n)
    offsets = D.offsets_of(D.CRITEO_CAT_SIZES, first=13)
    lines = []
    for i in range(n):
        num = [int(rng.integers(0, m + 1)) for m in D.CRITEO_MAX_VALS]
        cat = [int(rng.integers(0, s)) for s in D.CRITEO_CAT_SIZES]
        lines.append("%d\t%s\t%s" % (rng.integers(0, 2), "\t".join(map(str, num)), "\t".join(map(str, cat))))
    text = "\n".join(lines) + "\n"
    dt = DeviceText(text.encode())
    ids, wts, y, status = dt.parse_criteo(offsets, D.CRITEO_MAX_VALS)
    check_status(status, expect=40)
    of, ox, oy = T.criteo_block(lines, offsets, D.CRITEO_MAX_VALS.tolist())
    assert np.array_equal(ids.cpu().numpy(), of)
    assert np.array_equal(wts.cpu().numpy(), ox)                  # integers / max: the same double division, then one cast
    assert np.array_equal(y.cpu().numpy(), oy)
    # short, empty-field and non-numeric lines are flagged
    bad = DeviceText(("1\t" + "\t".join(["3"] * 10) + "\n" + lines[0].replace("\t", "\t\t", 1) + "\n" + lines[0].replace("\t", "\tz", 1) + "\n").encode())
    s = bad.parse_criteo(offsets, D.CRITEO_MAX_VALS)[3].cpu().numpy().tolist()
    assert s[0] == 11 and s[1] == -2 and s[2] == -2


def test_fnn_wnzh_trains_from_device_parsed_text(tmp_path):
    """train.fm.txt bytes -> DeviceText -> FNN_wnzh.train_step_device == train() on the host-parsed batch (same masks)"""
    from deepctr_b200 import data as D
    from deepctr_b200.ingest import DeviceText
    from deepctr_b200.theano_fnn import FNN_wnzh
    rng = np.random.default_rng(77)
    X, F, d, N = 500, 16, 5, 64
    sizes = np.full(F, X // F)
    sizes[-1] += X - sizes.sum()
    feat_field = np.repeat(np.arange(F), sizes)
    table = rng.uniform(-0.1, 0.1, (X, d))
    ids = D.make_field_batch(rng, 2 * N, sizes)
    y = rng.integers(0, 2, 2 * N)
    path = str(tmp_path / "train.fm.txt")
    D.write_fm_samples(path, ids, y)
    fm = dict(w_0=0.05, table=table, feat_field=feat_field)
    a, b = FNN_wnzh(fm, batch_size=N, hidden1=24, hidden2=12), FNN_wnzh(fm, batch_size=N, hidden1=24, hidden2=12)
    dt = DeviceText(path)
    feats, _, yd, status = dt.parse_sparse(F, skip=0, pad_id=-1, with_wts=False)
    assert dt.n_lines == 2 * N and (status == F).all()
    hf, hy = D.load_fm_samples(path)
    for s in range(0, 2 * N, N):
        la, pa = a.train(hf[s:s + N], hy[s:s + N])
        lb, pb = b.train_step_device(feats[s:s + N], yd[s:s + N])
        assert la == lb and np.array_equal(pa, pb)
    assert np.array_equal(a.get_table(), b.get_table())


# ---- f4: SNN first layer as an embedding bag ----------------------------------------------------------------------------
@pytest.mark.parametrize("N,F,H,X,act,with_w", [(1, 1, 1, 3, "none", False), (37, 16, 200, 5000, "sigmoid", True), (64, 7, 130, 90, "tanh", True),
                                                (1000, 16, 200, 133465, "sigmoid", False), (9, 20, 300, 50, "sigmoid", True)])
def test_bag_fwd(L, N, F, H, X, act, with_w):
    """dctr_bag_fwd: out = act(b0 + sum_f w * table[id]) -- x_dense of get_fi_h1_y (SNN_DAE.py:234-260); pad columns are 0"""
    import deepctr_oracle as O
    rng = np.random.default_rng(N + H)
    ld, ldo = (H + 3) // 4 * 4, (H + 3) // 4 * 4 + 4
    table = np.zeros((X, ld), np.float32)
    table[:, :H] = rng.normal(0, 0.3, (X, H))
    ids = rng.integers(0, X, (N, F)).astype(np.int32)
    ids[:, -1] = ids[:, 0]                                   # a feature twice in one sample counts twice
    wts = (rng.random((N, F)) < 0.8).astype(np.float32) if with_w else None
    b0 = rng.normal(0, 0.1, H).astype(np.float32)
    t_ids, t_tab, t_b0 = torch.tensor(ids).cuda(), torch.tensor(table).cuda(), torch.tensor(b0).cuda()
    t_w = torch.tensor(wts).cuda() if with_w else None
    out = torch.full((N, ldo), 7.0, device="cuda")
    L.call("dctr_bag_fwd", t_ids.data_ptr(), L.ptr(t_w), t_tab.data_ptr(), X, ld, N, F, H, t_b0.data_ptr(), L.ACT[act],
           out.data_ptr(), ldo, None)
    rows = table[ids, :H].astype(np.float64)
    pre = (rows * (wts[..., None] if with_w else 1.0)).sum(1) + b0
    ref = O.activate(act, pre) if act != "none" else pre
    o = out.cpu().numpy()
    close(o[:, :H], ref, 1e-5, 2e-6)
    assert (o[:, H:] == 0).all()


@pytest.mark.parametrize("gemm,use_graph,lambda1", [("simt", False, 0.0), ("tc", True, 0.01)])
def test_snn_train_steps(monkeypatch, gemm, use_graph, lambda1):
    """SNN fine-tuning steps (SNN_DAE.py:276-288) against OracleSNN: ragged samples, a feature that every sample has (a
    512-long duplicate segment), a feature twice inside one sample, injected dropout masks, then predict()"""
    monkeypatch.setenv("DCTR_GEMM", gemm)
    import deepctr_oracle as O
    from deepctr_b200.snn import SNN
    rng = np.random.default_rng(808)
    X, J, H0, H1, H2, N = 3000, 12, 200, 48, 20, 128
    ww0, bb0 = rng.normal(0, 0.2, (X, H0)), rng.normal(0, 0.1, H0)
    Ws = [rng.normal(0, 1, (H0, H1)) / np.sqrt(H0), rng.normal(0, 1, (H1, H2)) / np.sqrt(H1)]
    bs = [rng.normal(0, 0.1, H1), rng.normal(0, 0.1, H2)]
    f32 = lambda a: np.asarray(a, np.float32)
    m = SNN(x_dim=X, max_feats=J, batch_size=N, lr=0.05, dropout=0.5, hidden0=H0, hidden1=H1, hidden2=H2, lambda1=lambda1,
            weights=(f32(ww0), f32(bb0), f32(Ws[0]), f32(bs[0]), f32(Ws[1]), f32(bs[1])))
    m.use_graph = use_graph
    m.engine.set_layer(2, rng.uniform(-0.3, 0.3, (H2, 1)), np.array([0.05]))          # w3 = 0 would leave gx = 0
    w3, b3 = m.engine.get_layer(2)
    o = O.OracleSNN(f32(ww0), f32(bb0), [f32(Ws[0]), f32(Ws[1]), w3], [f32(bs[0]), f32(bs[1]), b3])
    for step in range(4):
        feats = np.full((N, J), -1, dtype=np.int64)
        for t in range(N):
            k = int(rng.integers(1, J))
            feats[t, :k] = rng.integers(0, X, k)
        feats[:, 0] = 11
        feats[5, 1] = feats[5, 2] = 77
        y = rng.integers(0, 2, N)
        r1, r2 = (rng.random((1, H1)) < 0.5).astype(np.float32), (rng.random((1, H2)) < 0.5).astype(np.float32)
        m.set_masks(r1, r2)
        loss, p = m.train(feats, y)
        oloss, op = o.train_step([list(r) for r in feats], y, 0.05, r1.astype(np.float64), r2.astype(np.float64), lambda1)
        if lambda1 == 0.0:
            close([loss], [oloss], 2e-5)
        close(p, op, 5e-5)
        tw0, tb0, w1, b1, w2, b2, w3n, b3n = m.get_weights()
        close(tw0, o.ww0, 5e-5, 5e-6)
        close(tb0, o.bb0, 5e-5, 5e-6)
        close(w1, o.Ws[0], 5e-5, 5e-6)
        close(w2, o.Ws[1], 5e-5, 5e-6)
        close(w3n, o.Ws[2].reshape(-1), 5e-5, 5e-6)
    close(m.predict(feats[:100]), o.predict([list(r) for r in feats[:100]]), 5e-5)


# ---- the fold without its weight gather (0 / 1 weights, the 0s on one pad row) -----------------------------------------
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("N,F,d,X", [(300, 16, 11, 400), (5000, 16, 11, 3000), (64, 5, 200, 40)])
def test_scatter_masked_equals_weighted(L, mode, N, F, d, X):
    """dctr_scatter_sgd_masked(skip_row = pad) == dctr_scatter_sgd with weights 1 / 0-on-the-pad-row, bit for bit (the pad
    row is all zero, as in FNN_wnzh / SNN), including a pad segment that spans many tiles"""
    rng = np.random.default_rng(N + d + mode)
    ld = (d + 3) // 4 * 4
    pad = X - 1
    ids = rng.integers(0, X - 1, (N, F)).astype(np.int32)
    ids[rng.random((N, F)) < 0.3] = pad
    ids[:, 0] = 3                                             # a hot row as well
    wts = (ids != pad).astype(np.float32)
    table = np.zeros((X, ld), np.float32)
    table[:pad, :d] = rng.normal(0, 0.1, (pad, d))
    g = rng.normal(0, 1, (N, F * d + 4)).astype(np.float32)
    M = N * F
    t_ids, t_w, t_g = torch.tensor(ids).cuda(), torch.tensor(wts).cuda(), torch.tensor(g).cuda()
    rows, pos = torch.empty(M, dtype=torch.int32, device="cuda"), torch.empty(M, dtype=torch.int32, device="cuda")
    sws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    fws = torch.zeros(L.lib().dctr_scatter_workspace(M, ld), dtype=torch.uint8, device="cuda")
    L.call("dctr_sort_rows", t_ids.data_ptr(), M, X, rows.data_ptr(), pos.data_ptr(), sws.data_ptr(), sws.numel(), None)
    ta, tb = torch.tensor(table).cuda(), torch.tensor(table).cuda()
    L.call("dctr_scatter_sgd", mode, rows.data_ptr(), pos.data_ptr(), M, F, d, t_g.data_ptr() + 4, g.shape[1], d, t_w.data_ptr(),
           0.05, 0.999, ta.data_ptr(), X, ld, fws.data_ptr(), fws.numel(), None)
    L.call("dctr_scatter_sgd_masked", mode, rows.data_ptr(), pos.data_ptr(), M, F, d, t_g.data_ptr() + 4, g.shape[1], d, pad,
           0.05, 0.999, tb.data_ptr(), X, ld, fws.data_ptr(), fws.numel(), None)
    torch.cuda.synchronize()
    assert torch.equal(ta, tb) and not torch.equal(ta, torch.tensor(table).cuda())
    assert bool((tb[pad] == 0).all())
