"""CPU tests of host-side logic that needs no device: optimizer dispatch (builf_optimizer, tf_util.py:15-29), the SNN slot
packing (get_fi_h1_y's feature lists, SNN_DAE.py:241-249) and the parser status check."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


class _FakeEngine:
    def __init__(self):
        self.calls = []

    def enable_adam(self, eps):
        self.calls.append(("adam", eps))

    def enable_ftrl(self):
        self.calls.append(("ftrl",))


def test_optimizer_dispatch_follows_builf_optimizer():
    from deepctr_b200.models import _optimizer
    e = _FakeEngine()
    lr, log = _optimizer(["adam", 1e-4, 1e-8, "sum"], e)
    assert lr == 1e-4 and e.calls == [("adam", 1e-8)] and "epsilon" in log
    e = _FakeEngine()
    lr, log = _optimizer(["ftrl", 0.01, "mean"], e)
    assert lr == 0.01 and e.calls == [("ftrl",)] and "ftrl" in log
    e = _FakeEngine()
    lr, log = _optimizer(["sgd", 1e-3, "sum"], e)             # anything else is plain gradient descent (tf_util.py:26-28)
    assert lr == 1e-3 and e.calls == []
    lr, _ = _optimizer(["whatever", 0.5, "sum"], e)
    assert lr == 0.5 and e.calls == []


def test_snn_slot_packing():
    from deepctr_b200.snn import SNN, init_weight

    class Fake:
        F, pad_row = 5, 99
    ids, wts = SNN.slots(Fake(), np.array([[3, -1, 7], [1, 2, -1]]))
    assert ids.tolist() == [[3, 99, 7, 99, 99], [1, 2, 99, 99, 99]] and wts.tolist() == [[1, 0, 1, 0, 0], [1, 1, 0, 0, 0]]
    ids, wts = SNN.slots(Fake(), [[3, 7], [1, 2, 4, 5, 6], []])
    assert ids.tolist() == [[3, 7, 99, 99, 99], [1, 2, 4, 5, 6], [99] * 5] and wts.sum(1).tolist() == [2, 5, 0]
    with pytest.raises(ValueError):
        SNN.slots(Fake(), [[1, 2, 3, 4, 5, 6]])
    with pytest.raises(ValueError):
        SNN.slots(Fake(), np.zeros((2, 6), dtype=np.int64))
    # dl_utils.init_weight (dl_utils.py:44-56): x4 for sigmoid, zero bias, bound sqrt(6 / (fan_in + fan_out))
    w, b = init_weight(np.random.RandomState(1), 30, 20, "sigmoid")
    w2, _ = init_weight(np.random.RandomState(1), 30, 20, "tanh")
    assert np.allclose(w, 4 * w2) and (b == 0).all() and np.abs(w2).max() <= np.sqrt(6.0 / 50)


def test_parser_status_check():
    from deepctr_b200.ingest import check_status
    ok = torch.tensor([16, 16, 3], dtype=torch.int32)
    assert check_status(ok).tolist() == [16, 16, 3]
    with pytest.raises(ValueError):
        check_status(torch.tensor([16, -2], dtype=torch.int32))
    with pytest.raises(ValueError):
        check_status(torch.tensor([40, 39], dtype=torch.int32), expect=40)
    with pytest.raises(ValueError):
        check_status(torch.tensor([16, 17], dtype=torch.int32), max_nnz=16)


def test_chunk_shuffle_reproduces_collect():
    """ingest.chunk_shuffle_perm == the order `collect` (ipinyou.py:11-20) hands out: chunks of `size` lines, each shuffled in
    place with np.random.shuffle -- same RandomState, same order, including the short last chunk"""
    from deepctr_b200.ingest import chunk_shuffle_perm, take_lines
    lines = ["line %d" % i for i in range(1037)]

    def collect(it, size, rng):                    # the reference function, RNG made explicit
        buf = []
        for _ in range(size):
            try:
                buf.append(next(it))
            except StopIteration:
                break
        rng.shuffle(buf)
        return buf
    rng = np.random.RandomState(99)
    it, ref = iter(lines), []
    while True:
        buf = collect(it, 250, rng)
        if not buf:
            break
        ref.extend(buf)
    perm = chunk_shuffle_perm(len(lines), 250, np.random.RandomState(99))
    assert [lines[i] for i in perm] == ref and sorted(perm.tolist()) == list(range(1037))
    t = torch.arange(1037 * 3).view(1037, 3)
    a, b = take_lines(perm, t, None)
    assert b is None and a[:, 0].tolist() == (perm * 3).tolist()


def test_early_stop_and_recalibration_equal_the_driver_logic():
    """metrics.early_stop / recalibrate against the statement-by-statement restatement of baseline.py:262-281, 369-370"""
    import deepctr_oracle as O
    from deepctr_b200.metrics import early_stop, recalibrate
    rng = np.random.default_rng(3)
    hits = 0
    for trial in range(300):
        n = int(rng.integers(0, 60))
        errs = (0.7 + 0.05 * rng.standard_normal(n) + 0.002 * np.arange(n) * rng.choice([-1, 1])).tolist()
        skip, smooth, stop = int(rng.integers(1, 4)), int(rng.integers(1, 12)), int(rng.integers(2, 12))
        for metric in ("auc", "rmse"):
            for step, least in ((100, 10), (5, 10)):
                a = early_stop(step, errs, metric, least, skip, smooth, stop)
                b = O.early_stop_literal(step, errs, metric, least, skip, smooth, stop)
                assert a == b, (trial, n, skip, smooth, stop, metric)
                hits += a
    assert hits > 50
    # baseline.py's own windows (skip 1, smooth 10, stop 10) on a record that turns down
    rec = [0.70 + 0.001 * i for i in range(30)] + [0.73 - 0.002 * i for i in range(12)]
    assert not early_stop(1000, rec[:30]) and early_stop(1000, rec)
    p = rng.random(1000)
    np.testing.assert_allclose(recalibrate(p, 0.025), O.recalibrate_literal(p, 0.025), rtol=1e-15)
    np.testing.assert_allclose(recalibrate(torch.tensor(p), 0.025).numpy(), O.recalibrate_literal(p, 0.025), rtol=1e-12)


def test_line_cursor_follows_the_reference_readers():
    """ingest.LineCursor(wrap=True) == get_batch_xy / get_batch_sparse_tensor (baseline.py:201-243) simulated on an in-memory
    'file' (end of file -> reopen and top up); wrap=False == consecutive full batches with the ragged tail dropped"""
    from deepctr_b200.ingest import LineCursor

    class Reader:                                  # the two reference functions with the file replaced by a list of line numbers
        def __init__(self, n):
            self.n, self.it = n, None

        def sparse(self, size):
            if self.it is None:
                self.it = iter(range(self.n))
            got = []
            for _ in range(size):
                try:
                    got.append(next(self.it))
                except StopIteration:
                    self.it = None
                    break
            return got

        def batch(self, size):
            got = self.sparse(size)
            if len(got) == size:
                return got
            return got + self.sparse(size - len(got))
    for n, size in [(10, 3), (10, 5), (10, 10), (7, 4), (3, 8), (100, 33)]:
        ref, cur = Reader(n), LineCursor(n, wrap=True)
        for _ in range(25):
            assert cur.next(size).tolist() == ref.batch(size), (n, size)
    cur = LineCursor(10, wrap=False)
    seen = [cur.next(4) for _ in range(7)]
    assert [None if s is None else s.tolist() for s in seen] == [[0, 1, 2, 3], [4, 5, 6, 7], None, [0, 1, 2, 3], [4, 5, 6, 7], None, [0, 1, 2, 3]]


def test_mask_prefetch_keeps_the_stream_and_the_step_plumbing_runs():
    """MaskStreams.take() hands out exactly the pairs draw() would have, however prefetch() calls are interleaved; and the
    host side of a step (_finish_step: upload, launch, post-launch prefetch, fetch) runs end to end on fakes"""
    from unittest import mock
    from deepctr_b200.theano_fnn import FNN_wnzh, MaskStreams
    a, b = MaskStreams(30, 10, 0.5), MaskStreams(30, 10, 0.5)
    rng = np.random.default_rng(0)
    for step in range(40):
        if rng.random() < 0.6:
            b.prefetch()
        if rng.random() < 0.3:
            b.prefetch()                          # a second call before the pair is taken draws nothing
        ra, rb = a.draw(), b.take()
        assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1])

    class Stage:
        def __init__(self):
            self.host = {"mask1": np.zeros((1, 32), np.float32), "mask2": np.zeros((1, 12), np.float32)}
            self.uploads = 0

        def upload(self):
            self.uploads += 1

    class Eng:
        out = torch.cat([torch.arange(10, dtype=torch.float32), torch.zeros(2)])      # pred[8] | loss | - | id status words (0 = clean)

    m = FNN_wnzh.__new__(FNN_wnzh)               # host plumbing only: no engine, no device
    m.N, m._stage, m.engine, m.mask_streams, m._next_masks = 8, Stage(), Eng(), MaskStreams(30, 10, 0.5), None
    m._out_pinned = torch.zeros(12)
    m._out_np = m._out_pinned.numpy()
    m._out_i32 = m._out_np.view(np.int32)
    m._run_step = lambda: None
    ref = MaskStreams(30, 10, 0.5)
    with mock.patch("torch.cuda.current_stream") as cs:
        for step in range(3):
            m._pre_step()
            want = ref.draw()
            assert np.array_equal(m._stage.host["mask1"][0, :30], want[0][0]) and np.array_equal(m._stage.host["mask2"][0, :10], want[1][0])
            loss, pred = m._finish_step()
            assert loss == 8.0 and pred.tolist() == list(range(8)) and m.mask_streams._ahead is not None
        assert cs.return_value.synchronize.call_count == 3 and m._stage.uploads == 3
    m.set_masks(np.ones(30), np.zeros(10))        # an injected pair does not consume the stream
    m._pre_step()
    assert m._stage.host["mask1"][0, :30].all() and not m._stage.host["mask2"][0, :10].any()
    m._pre_step()
    want = ref.draw()
    assert np.array_equal(m._stage.host["mask1"][0, :30], want[0][0])


def test_device_fed_step_plumbing_runs_on_fakes():
    """FNN_wnzh.train_step_device and the SNN mask hooks, host side only (CPU tensors stand in for the device blocks)"""
    from unittest import mock
    from deepctr_b200.snn import SNN
    from deepctr_b200.theano_fnn import FNN_wnzh, MaskStreams

    class Stage:
        def __init__(self):
            self.host = {"mask1": np.zeros((1, 32), np.float32), "mask2": np.zeros((1, 12), np.float32)}
            self.dev = {k: torch.zeros(v.shape) for k, v in self.host.items()}

    class Eng:
        out = torch.cat([torch.arange(10, dtype=torch.float32), torch.zeros(2)])      # pred[8] | loss | - | id status words (0 = clean)
        y = torch.zeros(8)

    for cls in (FNN_wnzh, SNN):
        m = cls.__new__(cls)
        m.N, m._stage, m.engine, m.mask_streams, m._next_masks = 8, Stage(), Eng(), MaskStreams(30, 10, 0.5), None
        m._out_pinned = torch.zeros(12)
        m._out_np = m._out_pinned.numpy()
        m._out_i32 = m._out_np.view(np.int32)
        m._run_step = lambda: None
        m._pre_step()
        m._post_launch()
        assert m.mask_streams._ahead is not None
        if cls is FNN_wnzh:
            m.feats = torch.zeros(8, 16, dtype=torch.int32)
            with mock.patch("torch.cuda.current_stream"):
                loss, pred = m.train_step_device(torch.ones(8, 16, dtype=torch.int32), torch.ones(8))
            assert loss == 8.0 and bool((m.feats == 1).all()) and bool((m.engine.y == 1).all())
            assert np.array_equal(m._stage.dev["mask1"].numpy(), m._stage.host["mask1"])


def test_pipelined_call_returns_each_result_one_call_late_on_fakes():
    """models._Base._train_async, host side only: CPU tensors stand in for the pinned and device blocks (pin_memory patched out, CUDA
    streams / events mocked).  Call t stages into pinned block t % 2, the step sees batch t in its device block, the result of step
    t comes back from call t + 1, flush_async() returns the last one, and models attached to a process group are refused."""
    from unittest import mock
    from deepctr_b200.models import _Staging
    from deepctr_b200.theano_fnn import FNN_wnzh
    N, J = 8, 16
    with mock.patch.object(torch.Tensor, "pin_memory", lambda self: self), mock.patch("torch.cuda.Stream"), mock.patch("torch.cuda.Event"), \
            mock.patch("torch.cuda.stream"), mock.patch("torch.cuda.current_stream"):
        m = FNN_wnzh.__new__(FNN_wnzh)
        m.N, m.J, m.mask_streams, m._next_masks = N, J, None, None
        m._stage = _Staging([("feats", np.int32, (N, J)), ("y", np.float32, (N,))], "cpu")

        class Eng:
            dev = "cpu"
            out = torch.zeros(N + 4)
        m.engine = Eng()
        m._out_pinned = torch.zeros(N + 4)
        m._out_np = m._out_pinned.numpy()
        m._out_i32 = m._out_np.view(np.int32)
        seen = []

        def run_step():                      # "the step": pred = 2 y, loss = sum of the feature ids, read from the DEVICE block
            y, feats = m._stage.dev["y"], m._stage.dev["feats"]
            seen.append(int(feats[0, 0]))
            m.engine.out[:N] = 2 * y
            m.engine.out[N] = float(feats.sum())
        m._run_step = run_step
        batches = [(np.full((N, J), t + 1, np.int64), np.arange(N, dtype=np.float64) + 10 * t) for t in range(5)]
        got = [m.train_async(*b) for b in batches]
        blocks = m._stage._blocks
        assert got[0] is None and seen == [1, 2, 3, 4, 5]
        assert blocks[0][1]["feats"][0, 0] == 5 and blocks[1][1]["feats"][0, 0] == 4          # block t % 2 holds batch t
        got = got[1:] + [m.flush_async()]
        assert m.flush_async() is None
        for t, (loss, pred) in enumerate(got):
            assert loss == (t + 1) * N * J and np.array_equal(pred, 2 * batches[t][1].astype(np.float32))
        # a bad id is reported by the call that collects that step: the status words travel in the step's own out block
        def bad_step():
            m.engine.out[:N + 1] = 0
            m.engine.out[N + 2:] = torch.tensor([2, 77], dtype=torch.int32).view(torch.float32)
        m._run_step = bad_step
        assert m.train_async(*batches[0]) is None
        m._run_step = run_step
        with pytest.raises(IndexError, match="77"):
            m.train_async(*batches[1])
        assert m.flush_async()[0] == 2 * N * J
        # a synchronous call while a pipelined step is in flight would overwrite that step's staging block: refused until it is collected
        assert m.train_async(*batches[2]) is None
        with pytest.raises(RuntimeError, match="flush_async"):
            m.train(*batches[3])
        with pytest.raises(RuntimeError, match="flush_async"):
            m.predict(batches[3][0])
        assert m.flush_async()[0] == 3 * N * J
        m._no_step_in_flight()
        m.dp = object()
        with pytest.raises(NotImplementedError):
            m.train_async(*batches[0])


def test_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (no GPU involved): ONE JSON line on stdout with the arm's metric / unit / config, `impl`,
    a cpu_baseline describing the run and an e2e block without copies; the reference's own script rides in cpu_baseline when
    oracle/_ref is there."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "FNN train samples/s" and d["unit"] == "samples/s" and d["higher_is_better"] is True
    assert d["config"]["workload"] == "fnn_ipinyou" and d["config"]["batch_per_gpu"] == 4096 and d["steps"] == 2
    assert d["value"] > 0 and d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    b0 = d["cpu_baseline"].get("reference_script_B0")
    if b0 is not None and "unavailable" not in b0:
        assert b0["kind"] == "reference" and 0 < b0["value"] < d["value"]


def test_bench_algorithmic_costs_are_the_survey_figures():
    """bench.call_cost = SURVEY.md 8(d)'s algorithmic bytes / flops per sample (iPinYou shape: F = 16, d = 11): gather-concat 1 540 B,
    scatter-SGD 2 176 B, FM forward 840 B + the s row, Theano-FNN MLP forward + data gradients; the roofline fractions are these
    figures over measured time, so they are pinned here."""
    import bench
    N, F, d = 4096, 16, 11
    gather = (0, 0, 0, 10 ** 6, 12, N, F, d, 0, 1, 0, 0, 0, 180, None)
    assert bench.call_cost("dctr_gather_concat_fwd", gather) == ("byte", N * 1540)
    ip = gather[:8] + (1,) + gather[9:]
    assert bench.call_cost("dctr_gather_concat_fwd", ip) == ("byte", N * (1540 + 4 * 120))            # + 4 P for the inner products
    scatter = (0, 0, 0, N * F, F, d)
    assert bench.call_cost("dctr_scatter_sgd", scatter) == ("byte", N * 2176)
    assert bench.call_cost("dctr_scatter_sgd_masked", scatter) == ("byte", N * 2176)
    assert bench.call_cost("dctr_fm_fwd", gather) == ("byte", N * (840 + 4 * d))
    unit, flops = bench.call_cost("dctr_mlp2tc_train", (N, 177, 300, 100))
    assert unit == "flop" and flops == 4.0 * N * (177 * 300 + 300 * 100 + 100)                        # 2 of the 3 GEMM families of a step
    assert bench.call_cost("dctr_gemm_tc", (2, 177, 300, N)) == ("flop", 2.0 * 177 * 300 * N)
    # the sort's figure: one histogram read of the keys + (key, position) read and written per pass; 20-bit ids sort in 2 passes of 10 bits
    assert bench.call_cost("dctr_sort_rows", (0, N * F, 10 ** 6)) == ("byte", N * F * (4 + 2 * 16))
    assert bench.call_cost("dctr_sort_rows", (0, 1 << 24, 2 * 10 ** 7)) == ("byte", (1 << 24) * (4 + 4 * 16))
