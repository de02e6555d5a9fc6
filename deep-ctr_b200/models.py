"""Host-side mirror of the reference's model interface (the drop-in surface).

The reference classes build a TensorFlow graph and are driven through `sess.run([ptmzr, loss, train_preds],
feed_dict)` / `model.eval_preds.eval(feed_dict)` (baseline.py:351,362-366,421).  The classes here keep the
constructor signatures, attribute names, feed keys and checkpoint format, and route every step to the
sm_100a kernels of libdctr.so through engine.py.  numpy in, numpy out; the host<->device copies happen
inside (pinned staging buffers), so a call is an end-to-end step.

Deviations, all forced by the environment or by reference defects:
  * random init uses numpy Generator(PCG64(seed)) per variable instead of TF's Philox stream (Q11);
    parity runs always inject weights through the pickle path of `_init_argv[-1]` (tf_util.py:41-49);
  * 'sgd', 'adam' and 'ftrl' (tf_util.py:15-29) all run on the device; Adam and FTRL are dense sweeps, as in TF;
  * FNN_OP_L3 sizes h1_w from the forward's true width F*d + d*d + 1 (Q2).
"""
import os
import pickle

import numpy as np
import torch

from . import _lib as L
from .engine import FMEngine, StepEngine


class Placeholder(object):
    """Stands in for a tf.placeholder: only used as a feed_dict key."""

    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return "<placeholder %s>" % self.name


class Fetch(object):
    """Stands in for a fetchable tensor: `fetch.eval(feed_dict=...)` like tf.Tensor.eval (baseline.py:366)."""

    def __init__(self, model, kind):
        self.model, self.kind = model, kind

    def eval(self, feed_dict=None):
        return self.model._fetch(self.kind, feed_dict or {})


class Session(object):
    """Minimal tf.Session look-alike so the reference loop body reads the same (baseline.py:328-351)."""

    def __init__(self, model):
        self.model = model

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def run(self, fetches, feed_dict=None):
        single = not isinstance(fetches, (list, tuple))
        fl = [fetches] if single else list(fetches)
        m = self.model
        out = {}
        if any(f is m.ptmzr for f in fl):
            loss, preds = m.train_step(feed_dict)
            out = {"ptmzr": None, "loss": loss, "train_preds": preds}
        res = [out[f.kind] if f.kind in out else f.eval(feed_dict) for f in fl]
        return res[0] if single else res


def init_var_map(_init_argv, variables):
    """tf_util.py:41-82: load `{name: ndarray}` from the pickle named by _init_argv[-1]; every missing key is
    drawn ('random': uniform[a,b) or normal(mean, std), one seed per random variable, in declaration order)
    or zero-filled."""
    path = _init_argv[-1]
    if path:
        with open(path, "rb") as f:
            var_map = pickle.load(f, encoding="latin1")
        log = "init model from: %s, " % path
    else:
        var_map, log = {}, "random init, "
    method = _init_argv[0]
    a, b, seeds = _init_argv[1:-1]
    if method == "normal":
        log += "init method: %s(mean=%g, stddev=%g), seeds: %s\n" % (method, a, b, str(seeds))
    else:
        log += "init method: %s(minval=%g, maxval=%g), seeds: %s\n" % (method, a, b, str(seeds))
    j = 0
    for key, shape, action in variables:
        if key in var_map:
            var_map[key] = np.asarray(var_map[key], dtype=np.float32).reshape(shape)
            continue
        if action == "random":
            rng = np.random.Generator(np.random.PCG64(seeds[j]))
            j += 1
            if method == "normal":
                var_map[key] = rng.normal(a, b, size=shape).astype(np.float32)
            else:
                var_map[key] = rng.uniform(a, b, size=shape).astype(np.float32)
        else:
            var_map[key] = np.zeros(shape, dtype=np.float32)
    return var_map, log


def _optimizer(_ptmzr_argv, engine):
    """builf_optimizer, tf_util.py:15-29: 'adam' -> AdamOptimizer(lr, epsilon), 'ftrl' -> FtrlOptimizer(lr) (TF defaults),
    anything else -> GradientDescentOptimizer(lr).  Switches the engine and returns (lr, log)."""
    name = _ptmzr_argv[0]
    if name == "adam":
        lr, eps = _ptmzr_argv[1:3]
        engine.enable_adam(eps=float(eps))
        return float(lr), "optimizer: %s, learning rate: %g, epsilon: %g" % (name, lr, eps)
    if name == "ftrl":
        engine.enable_ftrl()
    lr = _ptmzr_argv[1]
    return float(lr), "optimizer: %s, learning rate: %g" % (name, lr)


class _Staging(object):
    """One pinned host buffer and one device buffer holding every per-step input block back to back (16-byte aligned),
    so a step costs ONE host->device copy; `host[name]` are numpy views, `dev[name]` torch views of the same layout."""

    def __init__(self, blocks, device):
        offs, total = {}, 0
        for name, dtype, shape in blocks:
            nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
            offs[name] = (total, nbytes, dtype, tuple(shape))
            total += (nbytes + 15) // 16 * 16
        self.nbytes = total
        self._offs = offs
        self.pinned = torch.zeros(total, dtype=torch.uint8).pin_memory()
        self.device = torch.zeros(total, dtype=torch.uint8, device=device)
        tdt = {np.int32: torch.int32, np.float32: torch.float32}
        self.host, self.dev = self._host_views(self.pinned), {}
        for name, (o, nb, dtype, shape) in offs.items():
            self.dev[name] = self.device[o:o + nb].view(tdt[dtype]).view(*shape)

    def _host_views(self, pinned):
        return {name: pinned[o:o + nb].numpy().view(dtype).reshape(shape) for name, (o, nb, dtype, shape) in self._offs.items()}

    def upload(self):
        self.device.copy_(self.pinned, non_blocking=True)

    def enable_pipeline(self):
        """two-deep pipelining (_Base._train_async): a second pinned block, so the host can stage step t+1 while the upload of step t
        may still be in flight, and a device-side landing buffer the copy stream fills while the running step reads `device`."""
        if not hasattr(self, "landing"):
            second = torch.zeros(self.nbytes, dtype=torch.uint8).pin_memory()
            self._blocks = [(self.pinned, self.host), (second, self._host_views(second))]
            self.landing = torch.zeros_like(self.device)

    def select(self, k):
        """`pinned` / `host[...]` now name block k (every staging helper looks them up at call time)"""
        self.pinned, self.host = self._blocks[k]


class _Base(object):
    """Shared host logic: staging, CUDA-graph replay of the step, chunked inference."""

    use_graph = True

    def _setup_host(self, engine, batch_size, F, extra_blocks=(), with_ids=True):
        self.engine = engine
        self.N, self.F = batch_size, F
        N = batch_size
        blocks = ([("ids", np.int32, (N, F)), ("wts", np.float32, (N, F))] if with_ids else []) + [("y", np.float32, (N,))]
        self._stage = _Staging(blocks + list(extra_blocks), engine.dev)
        # the engine reads its inputs straight from the staging buffer and writes pred | loss into one output buffer
        if with_ids:
            engine.ids, engine.wts = self._stage.dev["ids"], self._stage.dev["wts"]
        engine.y = self._stage.dev["y"]
        # pred [N] | loss | - | id status {count, one offending id} (int32 words written by the kernels, dctr_set_id_status)
        engine.out = torch.zeros(N + 4, dtype=torch.float32, device=engine.dev)
        engine.pred, engine.loss = engine.out[:N], engine.out[N:N + 1]
        self._status_ptr = engine.out.data_ptr() + 4 * (N + 2)
        self._out_pinned = torch.zeros(N + 4, dtype=torch.float32).pin_memory()
        self._out_np = self._out_pinned.numpy()
        self._out_i32 = self._out_np.view(np.int32)
        self._graph = None
        self._warm = 0
        self.h2d_bytes = self._stage.nbytes
        self.d2h_bytes = (N + 4) * 4

    def session(self):
        return Session(self)

    def _bind_status(self):
        """the kernels launched next from this thread report out-of-range ids into this model's output block"""
        L.call("dctr_set_id_status", self._status_ptr)

    def _check_ids(self):
        """after the D2H of the output block: raise for ids outside the table -- TF's InvalidArgument (FM.py:57), the Theano
        script's KeyError (FNN_wnzh.py:95).  The offending ids were read as row 0 and never written."""
        bad = int(self._out_i32[self.N + 2])
        if bad:
            one = int(self._out_i32[self.N + 3])
            self.engine.out[self.N + 2:].zero_()
            raise IndexError("%d feature id(s) outside the embedding table in this batch (e.g. %d): ids must lie in [0, X_dim)" % (bad, one))

    # -- subclasses: _device_step() issues the kernels of one train step on the current stream
    def _run_step(self):
        self._bind_status()
        if not self.use_graph:
            self._device_step()
            return
        if self._graph is None:
            if self._warm < 2:          # eager warm-up steps before capture (attribute setup, allocator)
                self._device_step()
                self._warm += 1
                return
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            # thread_local: NCCL's watchdog thread may query events while a data-parallel step is being captured
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                self._device_step()
            self._graph = g
            # capture does not execute: fall through to the first replay
        self._graph.replay()

    def _finish_step(self):
        """upload the staged inputs (one copy), run the step, fetch pred | loss (one copy)."""
        self._stage.upload()
        self._run_step()
        self._out_pinned.copy_(self.engine.out, non_blocking=True)
        self._post_launch()            # host work that can overlap the step running on the device
        torch.cuda.current_stream().synchronize()
        self._check_ids()
        return float(self._out_np[self.N]), self._out_np[:self.N].copy()

    def _no_step_in_flight(self):
        """the synchronous calls reuse the staging blocks at once: a pipelined step that has not been collected would lose its inputs"""
        p = getattr(self, "_pipe", None)
        if p is not None and p["pending"] is not None:
            raise RuntimeError("a pipelined step is still in flight: collect it with flush_async() before a synchronous call")

    def _stage_ids(self, ids, wts, y):
        h = self._stage.host
        np.copyto(h["ids"], ids, casting="unsafe")
        np.copyto(h["wts"], wts, casting="unsafe")
        np.copyto(h["y"], y, casting="unsafe")

    def _stage_and_step(self, ids, wts, y):
        self._no_step_in_flight()
        self._stage_ids(ids, wts, y)
        self._pre_step()
        return self._finish_step()

    # ---- two-deep pipelined training: an ADDITION next to the synchronous calls (the reference's loop is synchronous, baseline.py:351,
    # FNN_wnzh.py:296-298).  Call t stages batch t into pinned block t % 2, queues its upload on a copy stream into a landing buffer,
    # queues [landing -> input block, the step, pred | loss -> pinned out block t % 2] on the compute stream and returns the result of
    # step t - 1: the host packs batch t + 1 and the bus carries it while the device runs step t.  Same kernels, same order, same
    # results as the synchronous call; flush_async() returns the last step's result.
    def _pipe_setup(self):
        self._stage.enable_pipeline()
        second = torch.zeros(self.N + 4, dtype=torch.float32).pin_memory()
        dev = self.engine.dev
        self._pipe = dict(k=0, pending=None, copy_stream=torch.cuda.Stream(device=dev), landing_free=None,
                          outs=[(self._out_pinned, self._out_np, self._out_i32), (second, second.numpy(), second.numpy().view(np.int32))],
                          h2d=[torch.cuda.Event(), torch.cuda.Event()], out=[torch.cuda.Event(), torch.cuda.Event()])
        return self._pipe

    def _pipe_result(self, k):
        p = self._pipe
        p["out"][k].synchronize()
        self._out_pinned, self._out_np, self._out_i32 = p["outs"][k]
        self._check_ids()
        return float(self._out_np[self.N]), self._out_np[:self.N].copy()

    def _train_async(self, stage_fn):
        if getattr(self, "dp", None) is not None or getattr(self, "sharded", None) is not None:
            raise NotImplementedError("the pipelined call has only been run on single-GPU models; data-parallel and row-sharded models "
                                      "(dist.attach / attach_sharded) take the synchronous call")
        p = getattr(self, "_pipe", None) or self._pipe_setup()
        st, k = self._stage, p["k"]
        # pinned block k was uploaded two calls ago; that step's result has been read since, so the upload is complete
        st.select(k)
        stage_fn()
        self._pre_step()
        main, cs = torch.cuda.current_stream(self.engine.dev), p["copy_stream"]
        if p["landing_free"] is not None:
            cs.wait_event(p["landing_free"])                 # the previous step has taken its inputs out of the landing buffer
        with torch.cuda.stream(cs):
            st.landing.copy_(st.pinned, non_blocking=True)
            p["h2d"][k].record(cs)
        main.wait_event(p["h2d"][k])
        st.device.copy_(st.landing, non_blocking=True)
        p["landing_free"] = torch.cuda.Event()
        p["landing_free"].record(main)
        # a later step is launched before an earlier one's id status has been looked at: every pipelined step starts from clean status
        # words, so the count that travels back with its result is its own
        self.engine.out[self.N + 2:].zero_()
        self._run_step()
        p["outs"][k][0].copy_(self.engine.out, non_blocking=True)
        p["out"][k].record(main)
        self._post_launch()
        prev, p["pending"], p["k"] = p["pending"], k, k ^ 1
        return None if prev is None else self._pipe_result(prev)

    def flush_async(self):
        """(loss, preds) of the last pipelined step, or None when nothing is pending."""
        p = getattr(self, "_pipe", None)
        if not p or p["pending"] is None:
            return None
        prev, p["pending"] = p["pending"], None
        return self._pipe_result(prev)

    def _fetch_out(self):
        self._out_pinned.copy_(self.engine.out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self._check_ids()
        return float(self._out_np[self.N]), self._out_np[:self.N].copy()

    def train_step_device(self, ids, wts, y):
        """One train step from DEVICE-resident blocks (ingest.DeviceText.parse_*): int32 global ids [N, F], float32 weights
        [N, F] (None = 1), float32 labels [N].  Nothing but pred | loss crosses the bus."""
        d = self._stage.dev
        d["ids"].copy_(ids)
        if wts is None:
            d["wts"].fill_(1.0)
        else:
            d["wts"].copy_(wts)
        d["y"].copy_(y)
        self._pre_step()
        self._run_step()
        self._post_launch()
        return self._fetch_out()

    def _predict_chunks(self, ids, wts, out_dev=None):
        """Inference over E rows in chunks of the static batch (last chunk padded with row 0 / weight 0).  With `out_dev`
        (a float32 device tensor of E elements) the predictions stay on the device and nothing is copied back."""
        self._no_step_in_flight()
        e, h = self.engine, self._stage.host
        E = ids.shape[0]
        out = np.empty(E, dtype=np.float32) if out_dev is None else None
        saved = self._swap_masks(None)
        self._bind_status()
        for s in range(0, E, self.N):
            k = min(self.N, E - s)
            h["ids"][:k] = ids[s:s + k]
            h["wts"][:k] = wts[s:s + k]
            if k < self.N:
                h["ids"][k:] = 0
                h["wts"][k:] = 0
            self._stage.upload()
            e.forward(False)
            if out_dev is not None:
                out_dev[s:s + k].copy_(e.pred[:k])
            self._out_pinned.copy_(e.out, non_blocking=True)
            torch.cuda.current_stream().synchronize()           # the staging buffer is reused by the next chunk
            self._check_ids()
            if out_dev is None:
                out[s:s + k] = self._out_np[:k]
        self._swap_masks(saved)
        return out if out_dev is None else out_dev

    def evaluate(self, ids, wts, labels, nds_rate=None):
        """Predict E rows (global ids [E, F], weights [E, F]) and score them ON THE DEVICE: dict(auc, log_loss, rmse, n_pos) --
        the eval half of the reference's loop (baseline.py:357-370 + watch_train :246-259) without the D2H of the predictions."""
        from .metrics import device_metrics
        ids, wts = np.asarray(ids).reshape(-1, self.F), np.asarray(wts).reshape(-1, self.F)
        preds = torch.empty(ids.shape[0], dtype=torch.float32, device=self.engine.dev)
        self._predict_chunks(ids, wts, out_dev=preds)
        return device_metrics(preds, np.asarray(labels, dtype=np.float32).reshape(-1), device=self.engine.dev, nds_rate=nds_rate)

    def _pre_step(self):
        pass

    def _post_launch(self):
        pass

    def _swap_masks(self, masks):
        return None


# ==================================================================================================
# FM  (FM.py)
# ==================================================================================================
class FM(_Base):
    """FM(batch_size, [X_dim, X_feas, rank], _init_argv, _ptmzr_argv, [lambda], mode, eval_size) -- FM.py:5."""

    def __init__(self, batch_size, _rch_argv, _init_argv, _ptmzr_argv, _reg_argv, mode, eval_size, device="cuda"):
        X_dim, X_feas, rank = _rch_argv
        self.X_dim, self.X_feas, self.rank, self.mode = X_dim, X_feas, rank, mode
        self.log = "input dim: %d, features: %d, rank: %d, " % (X_dim, X_feas, rank)
        var_map, log = init_var_map(_init_argv, [("W", [X_dim, 1], "random"), ("V", [X_dim, rank], "random"),
                                                 ("b", [1], "zero")])
        self.log += log
        eng = FMEngine(X_dim, X_feas, rank, batch_size, device=device)
        eng.set_table(np.concatenate([var_map["W"], var_map["V"]], axis=1))
        eng.bias().copy_(torch.as_tensor(var_map["b"], dtype=torch.float32))
        self._setup_host(eng, batch_size, X_feas)
        self._finish_init(_ptmzr_argv, _reg_argv, mode, eval_size)

    def _finish_init(self, _ptmzr_argv, _reg_argv, mode, eval_size):
        self.sp_id_hldr, self.sp_wt_hldr, self.lbl_hldr = Placeholder("sp_id"), Placeholder("sp_wt"), Placeholder("lbl")
        if mode == "train":
            self._lambda = float(_reg_argv[0])
            self._reduce = _ptmzr_argv[-1]
            self.lr, log = _optimizer(_ptmzr_argv, self.engine)
            self.log += "%s, lambda(l2): %g, reduce by: %s" % (log, self._lambda, self._reduce)
            self.eval_size = eval_size
            self.eval_id_hldr, self.eval_wts_hldr = Placeholder("eval_id"), Placeholder("eval_wts")
            self.eval_wt_hldr = self.eval_wts_hldr      # ipinyou.py:192 uses this spelling
            self.ptmzr, self.loss, self.train_preds = Fetch(self, "ptmzr"), Fetch(self, "loss"), Fetch(self, "train_preds")
            self.eval_preds = Fetch(self, "eval_preds")
        else:
            self.test_preds = Fetch(self, "test_preds")

    def _device_step(self):
        gs = 1.0 if self._reduce == "sum" else 1.0 / getattr(self, "global_N", self.N)
        self.engine.train_step(self.lr, self._lambda, gs)

    def train_step(self, feed):
        ids = np.asarray(feed[self.sp_id_hldr]).reshape(self.N, self.F)
        wts = np.asarray(feed[self.sp_wt_hldr]).reshape(self.N, self.F)
        y = np.asarray(feed[self.lbl_hldr]).reshape(self.N)
        return self._stage_and_step(ids, wts, y)

    def train_step_async(self, feed):
        """train_step, two-deep pipelined: returns (loss, preds) of the PREVIOUS call (None on the first); flush_async() ends the run."""
        ids = np.asarray(feed[self.sp_id_hldr]).reshape(self.N, self.F)
        wts = np.asarray(feed[self.sp_wt_hldr]).reshape(self.N, self.F)
        y = np.asarray(feed[self.lbl_hldr]).reshape(self.N)
        return self._train_async(lambda: self._stage_ids(ids, wts, y))

    def _fetch(self, kind, feed):
        if kind == "eval_preds":
            ids, wts = feed[self.eval_id_hldr], feed[self.eval_wts_hldr]
        elif kind == "test_preds":
            ids, wts = feed[self.sp_id_hldr], feed[self.sp_wt_hldr]
        else:
            raise ValueError("fetch %r needs Session.run with model.ptmzr" % kind)
        ids = np.asarray(ids).reshape(-1, self.F)
        return self._predict_chunks(ids, np.asarray(wts).reshape(-1, self.F))

    def dump(self, model_path):
        """FM.py:66-69: pickle {'W','V','b'}."""
        t = self.engine.get_table()
        var_map = {"W": t[:, :1].copy(), "V": t[:, 1:].copy(), "b": self.engine.bias().cpu().numpy()}
        with open(model_path, "wb") as f:
            pickle.dump(var_map, f)
        print("model dumped at %s" % model_path)


class LR(FM):
    """LR(batch_size, [X_dim, X_feas], _init_argv, _ptmzr_argv, [lambda], mode, eval_size) -- LR.py:4-58: the FM path with
    rank 0 (table rows hold the single weight `W`), same feeds; the eval weights placeholder is `eval_wt_hldr` (LR.py:44)."""

    def __init__(self, batch_size, _rch_argv, _init_argv, _ptmzr_argv, _reg_argv, mode, eval_size, device="cuda"):
        X_dim, X_feas = _rch_argv
        self.X_dim, self.X_feas, self.rank, self.mode = X_dim, X_feas, 0, mode
        self.log = "input dim: %d, features: %d, " % (X_dim, X_feas)
        var_map, log = init_var_map(_init_argv, [("W", [X_dim, 1], "random"), ("b", [1], "zero")])
        self.log += log
        eng = FMEngine(X_dim, X_feas, 0, batch_size, device=device)
        eng.set_table(var_map["W"])
        eng.bias().copy_(torch.as_tensor(var_map["b"], dtype=torch.float32))
        self._setup_host(eng, batch_size, X_feas)
        self._finish_init(_ptmzr_argv, _reg_argv, mode, eval_size)

    def dump(self, model_path):
        """LR.py:55-58: pickle {'W','b'}."""
        var_map = {"W": self.engine.get_table()[:, :1].copy(), "b": self.engine.bias().cpu().numpy()}
        with open(model_path, "wb") as f:
            pickle.dump(var_map, f)
        print("model dumped at %s" % model_path)


# ==================================================================================================
# FNN family  (FNN.py, FNN_IP_L3/L5/L7.py, FNN_OP_L3.py)
# ==================================================================================================
class _FNNBase(_Base):
    cross = "none"
    n_hidden = 2

    def __init__(self, cat_sizes, offsets, batch_size, _rch_argv, _init_argv, _ptmzr_argv, _reg_argv, mode,
                 eval_size, device="cuda", shard=None):
        """shard=(rank, world): row-sharded table (dist.attach_sharded) -- the engine then holds only the rows
        g with g % world == rank; without a checkpoint they are drawn on the device (one stream per shard)."""
        X_dim, X_feas, rank = _rch_argv[:3]
        hidden = list(_rch_argv[3:3 + self.n_hidden])
        act_func = _rch_argv[3 + self.n_hidden]
        self.X_dim, self.X_feas, self.rank, self.mode = X_dim, X_feas, rank, mode
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self.n_num = X_feas - len(cat_sizes)        # 13 numeric fields for Criteo (FNN.py:41-43)
        d = rank + 1
        crossw = {"none": 0, "ip": X_feas * (X_feas - 1) // 2, "op": d * d}[self.cross]
        mbd_dim = X_feas * d + crossw + 1
        self.log = "input dim: %d, features: %d, rank: %d, embedding: %d, " % (X_dim, X_feas, rank, mbd_dim)
        self.log += ", ".join("h%d: %d" % (i + 1, h) for i, h in enumerate(hidden)) + ", activate: %s" % act_func
        dims = [mbd_dim] + hidden + [1]
        device_table = shard is not None and not _init_argv[-1]
        variables = ([] if device_table else [("W", [X_dim, 1], "random"), ("V", [X_dim, rank], "random")]) + [("b", [1], "zero")]
        for i in range(len(dims) - 1):
            variables += [("h%d_w" % (i + 1), [dims[i], dims[i + 1]], "random"), ("h%d_b" % (i + 1), [dims[i + 1]], "zero")]
        var_map, log = init_var_map(_init_argv, variables)
        self.log += log
        act = act_func if act_func in ("tanh", "relu") else "sigmoid"     # tf_util.py:32-38
        self.shard = shard
        X_rows = X_dim if shard is None else (X_dim + shard[1] - 1) // shard[1]
        eng = StepEngine(X_rows, X_feas, rank, hidden, batch_size, cross=self.cross, bias_first=False,
                         act0=act, act=act, scatter_mode="sum", device=device)
        if device_table:
            gen = torch.Generator(device=eng.dev)
            gen.manual_seed(int(_init_argv[3][0]) * 1000003 + shard[0])
            a, b = float(_init_argv[1]), float(_init_argv[2])
            if _init_argv[0] == "normal":
                eng.table[:, :d].normal_(a, b, generator=gen)
            else:
                eng.table[:, :d].uniform_(a, b, generator=gen)
        else:
            full = np.concatenate([var_map["W"], var_map["V"]], axis=1)
            eng.set_table(full if shard is None else np.ascontiguousarray(full[shard[0]::shard[1]]))
        eng.bias().copy_(torch.as_tensor(var_map["b"], dtype=torch.float32))
        for i in range(len(dims) - 1):
            eng.set_layer(i, var_map["h%d_w" % (i + 1)], var_map["h%d_b" % (i + 1)])
        eng.wts_are_ones = False
        self._setup_host(eng, batch_size, X_feas)
        self.v_wt_hldr, self.c_id_hldr, self.c_wt_hldr = Placeholder("v_wt"), Placeholder("c_id"), Placeholder("c_wt")
        self.lbl_hldr = Placeholder("lbl")
        self._num_ids = np.arange(self.n_num, dtype=np.int64)[None, :]
        self._offsets64 = np.ascontiguousarray(self.offsets, dtype=np.int64)
        self._pack_threads = max(1, min(8, (os.cpu_count() or 2) // 2))
        self._masks = None
        if mode == "train":
            self._keep_prob = float(_reg_argv[0])
            self._reduce = _ptmzr_argv[-1]
            self.lr, log = _optimizer(_ptmzr_argv, eng)
            self.log += "%s, reduce by: %s, keep_prob(drop_out): %g" % (log, self._reduce, self._keep_prob)
            self.eval_size = eval_size
            self.eval_id_hldr, self.eval_wts_hldr = Placeholder("eval_id"), Placeholder("eval_wts")
            self.ptmzr, self.loss, self.train_preds = Fetch(self, "ptmzr"), Fetch(self, "loss"), Fetch(self, "train_preds")
            self.eval_preds = Fetch(self, "eval_preds")
            if self._keep_prob < 1.0:
                # tf.nn.dropout on every layer input incl. act(z1) (FNN.py:87-89): Bernoulli(keep) / keep.  All masks of a step
                # live in one buffer (layer 0 first) and are drawn by ONE kernel inside the captured step (dctr_dropout_masks,
                # Philox keyed by a device-side step counter), which also applies the first layer's mask to act(z1).
                sizes = [batch_size * eng.ldi[l] for l in range(eng.nl)]
                flat = torch.ones(sum(sizes), device=eng.dev)
                self._masks, o = [], 0
                for l, n in enumerate(sizes):
                    self._masks.append(flat[o:o + n].view(batch_size, eng.ldi[l]))
                    o += n
                eng.masks = list(self._masks)
                eng.dropout = (flat, self._keep_prob, 0x5EED, torch.zeros(2, dtype=torch.int32, device=eng.dev))
        else:
            self.test_preds = Fetch(self, "test_preds")

    def set_dropout_masks(self, masks):
        """Inject explicit masks (list over layers, [N, width] arrays or None) -- parity runs."""
        eng = self.engine
        self._masks = None
        eng.masks = [None] * eng.nl
        eng.dropout = None
        for l, m in enumerate(masks):
            if m is not None:
                t = torch.zeros(self.N, eng.ldi[l], device=eng.dev)
                t[:, :eng.widths[l]] = torch.as_tensor(np.asarray(m), dtype=torch.float32).to(eng.dev)
                eng.masks[l] = t
        self._graph = None

    def _swap_masks(self, masks):
        old = self.engine.masks
        self.engine.masks = masks if masks is not None else [None] * self.engine.nl
        return old

    def _device_step(self):
        gs = 1.0 if self._reduce == "sum" else 1.0 / getattr(self, "global_N", self.N)
        self.engine.train_step(self.lr, 1.0, gs, train_bias=True)

    def _global(self, v_wts, c_ids, c_wts):
        n = len(c_ids)
        ids = np.concatenate([np.broadcast_to(self._num_ids, (n, self.n_num)), np.asarray(c_ids) + self.offsets], axis=1)
        wts = np.concatenate([np.asarray(v_wts).reshape(n, self.n_num), np.asarray(c_wts)], axis=1)
        return ids, wts

    def _pack_feed(self, v_wts, c_ids, c_wts):
        """feed -> staging blocks in one native pass (dctr_host_pack_fields): the offset add of FNN.py:65-69, the int32
        cast and the numeric | categorical concatenation, without numpy temporaries."""
        h = self._stage.host
        N, n_cat = self.N, self.X_feas - self.n_num
        c_ids = np.ascontiguousarray(c_ids)
        if c_ids.dtype not in (np.int32, np.int64):
            c_ids = c_ids.astype(np.int64)
        v_wts = np.ascontiguousarray(v_wts, dtype=np.float32)
        c_wts = np.ascontiguousarray(c_wts, dtype=np.float32)
        if c_ids.shape != (N, n_cat) or c_wts.shape != (N, n_cat) or v_wts.size != N * self.n_num:
            raise ValueError("feed shapes %r %r %r do not match batch %d x (%d + %d) fields"
                             % (v_wts.shape, c_ids.shape, c_wts.shape, N, self.n_num, n_cat))
        L.call("dctr_host_pack_fields", c_ids.ctypes.data, c_ids.dtype.itemsize, self._offsets64.ctypes.data,
               v_wts.ctypes.data if self.n_num else None, c_wts.ctypes.data, N, self.n_num, n_cat,
               h["ids"].ctypes.data, h["wts"].ctypes.data, self._pack_threads)

    def _stage_feed(self, feed):
        self._pack_feed(feed[self.v_wt_hldr], feed[self.c_id_hldr], feed[self.c_wt_hldr])
        np.copyto(self._stage.host["y"], np.asarray(feed[self.lbl_hldr]).reshape(self.N), casting="unsafe")

    def train_step(self, feed):
        self._no_step_in_flight()
        self._stage_feed(feed)
        self._pre_step()
        return self._finish_step()

    def train_step_async(self, feed):
        """train_step, two-deep pipelined: the pack and the upload of this batch run under the previous step.  Returns (loss, preds)
        of the PREVIOUS call (None on the first); flush_async() returns the last one."""
        return self._train_async(lambda: self._stage_feed(feed))

    def _fetch(self, kind, feed):
        if kind == "eval_preds":       # global ids [E, X_feas]; the graph subtracts offsets itself (FNN.py:65-69)
            ids, wts = np.asarray(feed[self.eval_id_hldr]), np.asarray(feed[self.eval_wts_hldr])
        elif kind == "test_preds":
            ids, wts = self._global(feed[self.v_wt_hldr], feed[self.c_id_hldr], feed[self.c_wt_hldr])
        else:
            raise ValueError("fetch %r needs Session.run with model.ptmzr" % kind)
        return self._predict_chunks(ids, wts)

    def _global_table(self):
        """[X_dim, d] table in global row order.  Row-sharded mode (shard = (rank, world)): the engine holds the rows g with
        g % world == rank as local rows g // world; the shards are all-gathered (a collective: every rank must call) and
        interleaved back, so the checkpoint is the reference's [X_dim, ...] format whatever the sharding."""
        t = self.engine.get_table()
        if self.shard is None:
            return t
        rank, world = self.shard
        sh = getattr(self, "sharded", None)
        if sh is None or world == 1:
            return t[:self.X_dim]
        loc = self.engine.table[:, :self.engine.d].contiguous()
        parts = [torch.empty_like(loc) for _ in range(world)]
        sh.dist.all_gather(parts, loc, group=sh.group)
        full = np.empty((self.X_dim, self.engine.d), dtype=np.float32)
        for r, part in enumerate(parts):
            rows = full[r::world]
            rows[...] = part[:rows.shape[0]].cpu().numpy()
        return full

    def var_map(self):
        t = self._global_table()
        vm = {"W": t[:, :1].copy(), "V": t[:, 1:].copy(), "b": self.engine.bias().cpu().numpy()}
        for i in range(self.engine.nl):
            w, b = self.engine.get_layer(i)
            vm["h%d_w" % (i + 1)], vm["h%d_b" % (i + 1)] = w, b
        return vm

    def dump(self, model_path):
        """FNN.py:96-101: pickle {'W','V','b','h{i}_w','h{i}_b'}.  Row-sharded: a collective call, rank 0 writes the file."""
        vm = self.var_map()
        if self.shard is not None and self.shard[0] != 0:
            return
        with open(model_path, "wb") as f:
            pickle.dump(vm, f)
        print("model dumped at %s" % model_path)


class FNN(_FNNBase):
    """FNN.py:4 -- _rch_argv = [X_dim, X_feas, rank, h1, h2, act]."""
    cross, n_hidden = "none", 2


class FNN_IP_L3(_FNNBase):
    """FNN_IP_L3.py:4 -- _rch_argv = [X_dim, X_feas, rank, h1, h2, h3, act]."""
    cross, n_hidden = "ip", 3


class FNN_IP_L5(_FNNBase):
    """FNN_IP_L5.py:4 -- five hidden layers."""
    cross, n_hidden = "ip", 5


class FNN_IP_L7(_FNNBase):
    """FNN_IP_L7.py:4 -- seven hidden layers."""
    cross, n_hidden = "ip", 7


class FNN_OP_L3(_FNNBase):
    """FNN_OP_L3.py:4 -- outer-product layer; h1_w sized from the true forward width."""
    cross, n_hidden = "op", 3
