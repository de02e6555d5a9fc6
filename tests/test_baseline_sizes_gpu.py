"""Parity at the sizes BASELINE.json quotes -- the CUDA path through the C-ABI against the float64 oracle (itself checked
against the reference's own code in tests/test_oracle_vs_ref.py):
  * sort + fold (SUM, SEQ_DECAY; masked and weighted) on M = 2^20 slots over a 2 * 10^7-row table: 25-bit keys (the 8-bit,
    4-pass sort), one row with 70 000 occurrences and one with 20 000 (the Criteo numeric rows of FNN.py:77 are hit by EVERY
    sample: at N = 65 536 their segments span 256+ fold tiles and several round trips of the stitch kernel);
  * one full train step of fnn_ipinyou and fm_ipinyou at N = 4 096, X = 10^6 (BASELINE configs 1/2) and of fnn_criteo at
    N = 65 536, X = 1 636 608 on the tensor-core GEMM path (config 5's shape on one GPU), bench.py's own model builders.
Tolerance: 1e-5 relative per step (north_star) with an absolute floor that scales with the length of the sums involved; the
measured maximum errors are printed."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RTOL = 1e-5


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def relerr(a, b, floor):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / (np.abs(b) + floor)))


def dev(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda()


@pytest.mark.parametrize("mode", ["sum", "seq_decay"])
@pytest.mark.parametrize("variant", ["masked", "weighted"])
def test_sort_and_fold_at_scale(mode, variant):
    import deepctr_oracle as O
    from deepctr_b200 import _lib as L
    rng = np.random.default_rng(77)
    N, F, d, ld, X = 65536, 16, 11, 12, 20_000_000
    M = N * F
    from deepctr_b200 import data as D
    sizes = D.ipinyou_field_sizes(X - 1)
    ids = D.make_field_batch(rng, N, sizes)                 # Zipf(1.1) within 16 fields
    hot_a, hot_b, pad = int(ids[0, 3]), int(ids[1, 7]), X - 1
    ids[:, 3] = np.where(rng.random(N) < 0.55, hot_a, ids[:, 3])        # ~36 000 occurrences in field 3 ...
    ids[:, 4] = np.where(rng.random(N) < 0.55, hot_a, ids[:, 4])        # ... plus ~36 000 in field 4: one row, ~72 000 occurrences
    ids[:, 7] = np.where(rng.random(N) < 0.30, hot_b, ids[:, 7])        # ~20 000
    ids[:, 15] = np.where(rng.random(N) < 0.10, pad, ids[:, 15])        # a pad row (weight 0)
    counts = np.bincount(ids.reshape(-1), minlength=1)
    assert counts[hot_a] >= 65536 and 18000 <= counts[hot_b] <= 22000
    wts = np.ones((N, F), np.float32)
    if variant == "weighted":
        wts = rng.uniform(0.5, 1.5, (N, F)).astype(np.float32)
    wts[ids == pad] = 0
    D0 = 1 + F * d
    g = (rng.standard_normal((N, D0)) * 1e-2).astype(np.float32)
    touched = np.unique(ids)
    rows0 = rng.uniform(-0.05, 0.05, (len(touched), d)).astype(np.float32)
    lr, decay = 0.01, 1 - 2 * 0.1 * 0.01 / N
    # ---- device ----
    table = torch.zeros(X, ld, device="cuda")
    table[dev(touched, torch.int64), :d] = dev(rows0, torch.float32)
    ids_d, wts_d, g_d = dev(ids, torch.int32), dev(wts, torch.float32), dev(g, torch.float32)
    rows_s, pos_s = torch.zeros(M, dtype=torch.int32, device="cuda"), torch.zeros(M, dtype=torch.int32, device="cuda")
    ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    sw = torch.zeros(L.lib().dctr_scatter_workspace(M, ld), dtype=torch.uint8, device="cuda")
    L.call("dctr_sort_rows", ids_d.data_ptr(), M, X, rows_s.data_ptr(), pos_s.data_ptr(), ws.data_ptr(), ws.numel(), None)
    flat = ids.reshape(-1)
    order = np.argsort(flat, kind="stable")
    assert np.array_equal(rows_s.cpu().numpy(), flat[order]) and np.array_equal(pos_s.cpu().numpy(), order)     # bit-exact, stable
    md = L.SCATTER_SEQ_DECAY if mode == "seq_decay" else L.SCATTER_SUM
    dk = decay if mode == "seq_decay" else 1.0
    if variant == "masked":
        L.call("dctr_scatter_sgd_masked", md, rows_s.data_ptr(), pos_s.data_ptr(), M, F, d, g_d.data_ptr() + 4, D0, d, pad, lr, dk,
               table.data_ptr(), X, ld, sw.data_ptr(), sw.numel(), None)
    else:
        L.call("dctr_scatter_sgd", md, rows_s.data_ptr(), pos_s.data_ptr(), M, F, d, g_d.data_ptr() + 4, D0, d, wts_d.data_ptr(), lr, dk,
               table.data_ptr(), X, ld, sw.data_ptr(), sw.numel(), None)
    torch.cuda.synchronize()
    got = table[dev(touched, torch.int64), :d].cpu().numpy()
    # ---- oracle (float64): FNN_wnzh.py:300-306 per occurrence / the summed TF gradient ----
    ge = g.astype(np.float64)[:, 1:].reshape(N, F, d) * wts.astype(np.float64)[..., None]
    want = rows0.astype(np.float64)
    compact = np.searchsorted(touched, ids)
    if mode == "seq_decay":
        if variant == "masked":       # the pad row's segment is never written
            keep = ids != pad
            O.seq_decay_scatter_closed_form(want, compact[keep], ge[keep], lr, decay)
        else:
            O.seq_decay_scatter_closed_form(want, compact, ge, lr, decay)
    else:
        acc = np.zeros_like(want)
        np.add.at(acc, compact.reshape(-1), ge.reshape(M, d))
        want -= lr * acc
        if variant == "masked":
            want[np.searchsorted(touched, pad)] = rows0[np.searchsorted(touched, pad)]
    # sums of n terms of size ~lr*|g|: fp32 error ~ sqrt(n) * eps * lr * |g| -- the floor of the comparison
    seg = counts[touched].astype(np.float64)
    floor = 1e-7 + np.sqrt(seg)[:, None] * 6e-8 * lr * 3e-2
    err = np.max(np.abs(got - want) / (np.abs(want) + floor / RTOL))
    print("fold %s/%s at M=2^20: max error %.2e of tolerance (longest segment %d)" % (mode, variant, err / RTOL, counts.max()))
    assert err <= RTOL


def _bench():
    sys.path.insert(0, ROOT)
    import bench
    return bench


def test_fnn_ipinyou_step_at_baseline_size():
    """BASELINE config 1/2 shape: FNN 177-300-100-1 in the Theano form, N = 4 096, X = 10^6, dropout masks drawn, sequential
    decayed scatter -- two steps (the second replays the captured CUDA graph path of the default bench)."""
    import deepctr_oracle as O
    bench = _bench()
    N = 4096
    torch.manual_seed(20260)             # the model's initial weights come from torch's generator: same numbers in every run
    m, train = bench.build_model("fnn_ipinyou", N)
    pool, sizes = bench.make_pool("theano", 16, 1_000_000, N, 3, 5)
    e = m.engine
    Ws = [e.get_layer(l)[0].astype(np.float64) for l in range(3)]
    bs = [e.get_layer(l)[1].astype(np.float64) for l in range(3)]
    table = np.concatenate([m.get_table().astype(np.float64), np.zeros((1, 11))], 0)
    net = O.OracleNet(table, float(e.bias().item()), Ws, bs, 16, 11, "none", True, "none", "tanh")
    rng = np.random.default_rng(1)
    for step, (ids, wts, y) in enumerate(pool):
        r1, r2 = rng.binomial(1, 0.5, (1, 300)).astype(np.float64), rng.binomial(1, 0.5, (1, 100)).astype(np.float64)
        m.set_masks(r1, r2)
        loss, p = m.train(ids, y)
        # magnitude of the terms a row's update is made of (the weekday / hour rows are hit by hundreds of samples of the batch:
        # the fp32 error of such a sum scales with sum |lr g|, not with the smaller |sum| the terms cancel to)
        logit, cache = net.forward(ids, None, [None, r1, r2])
        ge = net.backward(cache, O.sigmoid(logit) - y)[3]
        mag = np.zeros((net.table.shape[0], 1))
        np.add.at(mag, ids.reshape(-1), np.abs(ge).reshape(-1, 11).max(1, keepdims=True))
        # the same for the dense weights: a weight's update is lr * sum_n x_n dz_n over the 4 096 samples of the batch
        dpre, wmag = (O.sigmoid(logit) - y).reshape(-1, 1), [None] * 3
        for l in (2, 1, 0):
            a_in, mk = cache["ins"][l], cache["masks"][l]
            wmag[l] = np.abs(a_in).T @ np.abs(dpre)
            din = dpre @ net.Ws[l].T
            if mk is not None:
                safe = np.where(mk != 0, mk, 1.0)
                dpre = din * mk * O.activate_grad_from_out(net.act if l > 0 else net.act0, np.where(mk != 0, a_in / safe, 0.0))
            else:
                dpre = din * O.activate_grad_from_out(net.act if l > 0 else net.act0, a_in)
        ol, op = net.train_step_theano(ids, y, m.lr, m.lambda_fm, [None, r1, r2], m.lambda1, closed_form=True)
        tol = (step + 1) * RTOL
        e_p, e_l = relerr(p, op, 1e-3), abs(loss - ol) / abs(ol)
        e_t = float(np.max(np.abs(m.get_table() - net.table[:-1]) / (np.abs(net.table[:-1]) + m.lr * mag[:-1] + 1e-3)))
        e_w = max(float(np.max(np.abs(e.get_layer(l)[0] - net.Ws[l]) / (np.abs(net.Ws[l]) + m.lr * wmag[l].reshape(net.Ws[l].shape) + 1e-2)))
                  for l in range(3))
        print("fnn_ipinyou step %d (%s): loss %.2e preds %.2e table %.2e weights %.2e of 1" % (step, e.gemm_fn, e_l, e_p, e_t, e_w))
        assert max(e_p, e_l, e_t, e_w) <= tol


def test_fm_ipinyou_step_at_baseline_size():
    import deepctr_oracle as O
    bench = _bench()
    N = 4096
    m, train = bench.build_model("fm_ipinyou", N)
    pool, sizes = bench.make_pool("fm", 16, 1_000_000, N, 3, 6)
    ofm = O.OracleFM(m.engine.get_table().astype(np.float64), float(m.engine.bias().item()))
    rng = np.random.default_rng(2)
    for step, (ids, wts, y) in enumerate(pool):
        wts = rng.uniform(0.5, 1.5, wts.shape).astype(np.float32)
        loss, p = train(ids, wts, y)
        ol, op = ofm.train_step(ids, wts, y, 1e-3, 0.0)
        e_p, e_l = relerr(p, op, 1e-3), abs(loss - ol) / abs(ol)
        e_t = relerr(m.engine.get_table(), ofm.table, 1e-3)
        print("fm_ipinyou step %d: loss %.2e preds %.2e table %.2e" % (step, e_l, e_p, e_t))
        assert max(e_p, e_l, e_t) <= (step + 1) * RTOL


def test_fnn_criteo_step_at_baseline_size():
    """N = 65 536 samples x 39 fields over the reference's 1 636 608 rows, FNN 430-400-400-1: the tcgen05 GEMM inside a model
    step, the numeric rows 0..12 as 65 536-long fold segments, 2.5 M sort keys."""
    import deepctr_oracle as O
    bench = _bench()
    N = 65536
    m, train = bench.build_model("fnn_criteo", N)
    assert m.engine.gemm_fn == "dctr_gemm_tc"
    pool, sizes = bench.make_pool("tf", 39, 1_636_608, N, 2, 8)
    e = m.engine
    nl = e.nl
    vm = m.var_map()
    net = O.OracleNet(np.concatenate([vm["W"], vm["V"]], 1), float(vm["b"][0]), [vm["h%d_w" % (i + 1)] for i in range(nl)],
                      [vm["h%d_b" % (i + 1)] for i in range(nl)], 39, 11, "none", False, "tanh", "tanh")
    for step, (ids, wts, y) in enumerate(pool):
        loss, p = train(ids, wts, y)
        # magnitude of the terms each row's gradient sum is made of: the error of a 65 536-term fp32 sum (rows 0..12 are hit by
        # every sample) scales with sum |g|, not with the much smaller |sum g| the terms cancel to
        logit, cache = net.forward(ids, wts.astype(np.float64))
        ge = net.backward(cache, O.sigmoid(logit) - y)[3]
        mag = np.zeros((net.table.shape[0], 1))
        np.add.at(mag, ids.reshape(-1), np.abs(ge).reshape(-1, 11).max(1, keepdims=True))
        ol, op = net.train_step_tf(ids, wts.astype(np.float64), y, 1e-3)
        vm = m.var_map()
        e_p, e_l = relerr(p, op, 1e-3), abs(loss - ol) / abs(ol)
        touched = np.unique(ids)
        got_t = np.concatenate([vm["W"], vm["V"]], 1)[touched].astype(np.float64)
        e_t = float(np.max(np.abs(got_t - net.table[touched]) / (np.abs(net.table[touched]) + 1e-3 * mag[touched] + 1e-3)))
        # a weight gradient here is a sum over 65 536 samples: compare against the scale of the layer
        e_w = max(float(np.max(np.abs(vm["h%d_w" % (i + 1)] - net.Ws[i])) / np.max(np.abs(net.Ws[i]))) for i in range(nl))
        print("fnn_criteo step %d: loss %.2e preds %.2e touched rows %.2e weights %.2e (of max|W|)" % (step, e_l, e_p, e_t, e_w))
        assert max(e_p, e_l, e_w, e_t) <= (step + 1) * RTOL
