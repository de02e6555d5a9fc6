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
