"""Two-deep pipelined training calls (models._Base._train_async: FNN*.train_step_async, FM.train_step_async, FNN_wnzh.train_async) against
the synchronous calls of a second, identically initialised model: same kernels in the same order, so losses, predictions and the
updated parameters must be EQUAL bit for bit, one call late; eager steps and CUDA-graph replay; an out-of-range id surfaces as
IndexError when that step's result is collected."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _run_both(sync_call, async_call, flush, batches):
    want = [sync_call(*b) for b in batches]
    got = [async_call(*b) for b in batches]
    assert got[0] is None
    got = got[1:] + [flush()]
    assert flush() is None
    for (l0, p0), (l1, p1) in zip(want, got):
        assert l0 == l1 and np.array_equal(p0, p1)


@pytest.mark.parametrize("use_graph", [False, True])
def test_fnn_family_async_equals_sync(tmp_path, use_graph):
    import test_models_gpu as T
    N = 96
    models = []
    for _ in range(2):
        m, _, (cat_sizes, offsets, n_num) = T.make_fnn("FNN_IP_L3", tmp_path, np.random.default_rng(7), N, [40, 30, 20], "tanh", use_graph=use_graph)
        models.append(m)
    a, b = models
    rng = np.random.default_rng(8)
    batches = []
    for _ in range(6):
        v_wts, c_ids, c_wts, y, _, _ = T.fnn_batch(rng, N, cat_sizes, offsets, n_num)
        batches.append(({a.v_wt_hldr: v_wts, a.c_id_hldr: c_ids, a.c_wt_hldr: c_wts, a.lbl_hldr: y},))
    feed_b = lambda f: {b.v_wt_hldr: f[a.v_wt_hldr], b.c_id_hldr: f[a.c_id_hldr], b.c_wt_hldr: f[a.c_wt_hldr], b.lbl_hldr: f[a.lbl_hldr]}
    _run_both(a.train_step, lambda f: b.train_step_async(feed_b(f)), b.flush_async, batches)
    va, vb = a.var_map(), b.var_map()
    assert all(np.array_equal(va[k], vb[k]) for k in va)
    # the synchronous call still works on a model that has been pipelined (either pinned block may be current)
    la, pa = a.train_step(batches[0][0])
    lb, pb = b.train_step(feed_b(batches[0][0]))
    assert la == lb and np.array_equal(pa, pb)


@pytest.mark.parametrize("use_graph", [False, True])
def test_fm_async_equals_sync(use_graph):
    from deepctr_b200.models import FM
    N, X, F, rank = 128, 3000, 16, 10
    a, b = (FM(N, [X, F, rank], ["uniform", -0.05, 0.05, [0x3210, 0x7654], None], ["sgd", 0.05, "sum"], [1e-3], "train", 100) for _ in range(2))
    a.use_graph = b.use_graph = use_graph
    rng = np.random.default_rng(9)
    batches = [(rng.integers(0, X, (N, F)), rng.uniform(0.5, 1.5, (N, F)).astype(np.float32), rng.integers(0, 2, N)) for _ in range(6)]
    feed = lambda m, ids, wts, y: {m.sp_id_hldr: ids.reshape(-1), m.sp_wt_hldr: wts.reshape(-1), m.lbl_hldr: y}
    _run_both(lambda *t: a.train_step(feed(a, *t)), lambda *t: b.train_step_async(feed(b, *t)), b.flush_async, batches)
    assert np.array_equal(a.engine.get_table(), b.engine.get_table())


@pytest.mark.parametrize("use_graph,dropout", [(False, 0.5), (True, 0.5), (True, 1.0)])
def test_theano_fnn_async_equals_sync(use_graph, dropout):
    import test_models_gpu as T
    N = 100
    (a, _, X, F), (b, _, _, _) = (T.make_theano(np.random.default_rng(5), N, dropout=dropout, use_graph=use_graph) for _ in range(2))
    rng = np.random.default_rng(6)
    batches = [T.theano_batch(rng, N, X, F) for _ in range(6)]           # both models draw the same mask stream, in call order
    _run_both(a.train, b.train_async, b.flush_async, batches)
    assert np.array_equal(a.get_table(), b.get_table())
    assert all(np.array_equal(x, y) for x, y in zip(a.get_weights()[:5], b.get_weights()[:5]))


def test_async_reports_bad_ids_with_the_step_they_belong_to():
    import test_models_gpu as T
    N = 64
    m, _, X, F = T.make_theano(np.random.default_rng(5), N)
    rng = np.random.default_rng(6)
    good = T.theano_batch(rng, N, X, F)
    bad = (good[0].copy(), good[1])
    bad[0][3, 2] = X + 5
    assert m.train_async(*good) is None
    assert m.train_async(*bad) is not None          # returns the result of the first (good) step
    with pytest.raises(IndexError):
        m.train_async(*good)                        # collects the bad step
    m.flush_async()
