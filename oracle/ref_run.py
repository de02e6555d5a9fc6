"""Runs the reference's own Python (oracle/_ref, built by oracle/build_ref.py) on the engine shims of oracle/shims/ --
TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/ and tests/golden/make_golden_ref.py import this.

  run_fnn_wnzh(workdir)        executes the converted FNN_wnzh.py top to bottom (runpy) in <workdir>/python with the data
                               files it opens under <workdir>/data; returns its final globals and the log of every
                               train() / predict() call the script made (inputs, dropout draws, outputs).
  run_snn_dae(workdir, w)      the same for the converted SNN_DAE.py, started from the weight pickle it loads.
  tf_modules()                 imports tf_util / FM / LR / FNN / FNN_IP_L3 / L5 / L7 / FNN_OP_L3 on the tensorflow shim.
  load_baseline(workdir, ...)  imports baseline.py (its module body opens ../log/<tag>) so that get_fxy, get_batch_xy,
                               early_stop, watch_train and train() can be called with overridden module globals.
  load_ipinyou()               imports ipinyou.py (collect, stat, load_ipinyou_data).
"""
import contextlib
import importlib
import io
import linecache
import os
import runpy
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref", "python")
SHIMS = os.path.join(HERE, "shims")


def available():
    if not os.path.isdir(REF):
        sys.path.insert(0, HERE)
        try:
            import build_ref
            build_ref.build(verbose=False)
        finally:
            sys.path.remove(HERE)
    return os.path.isdir(REF)


@contextlib.contextmanager
def _paths(cwd=None):
    """reference dir + shims in front of sys.path (the reference imports its siblings by bare name), optional chdir."""
    old_cwd = os.getcwd()
    sys.path[:0] = [REF, SHIMS]
    try:
        if cwd:
            os.chdir(cwd)
        yield
    finally:
        os.chdir(old_cwd)
        for p in (REF, SHIMS):
            sys.path.remove(p)


def _fresh(*names):
    for n in names:
        sys.modules.pop(n, None)


def run_fnn_wnzh(workdir, argv=(), quiet=True, budget_s=None):
    """workdir/data/{fm.model.txt, train.fm.txt, test.fm.txt} must exist.  Returns (globals, calls).
    budget_s: stop the script at its first train() / predict() call more than that many seconds after its first one (the shim raises
    BudgetExceeded, caught here); the globals are then only {'__stdout__', '__stopped__'} and `calls` holds what ran -- bench.py's bounded timing sample."""
    assert available(), "oracle/_ref is not built (needs /root/reference)"
    py = os.path.join(workdir, "python")
    os.makedirs(py, exist_ok=True)
    os.makedirs(os.path.join(workdir, "log"), exist_ok=True)
    _fresh("theano", "theano.tensor", "theano.tensor.shared_randomstreams", "dl_utils")
    linecache.clearcache()          # the script reads '../data/train.fm.txt' through linecache: a relative key, stale across runs
    old_argv = sys.argv
    sys.argv = ["FNN_wnzh.py"] + list(argv)
    buf = io.StringIO()
    try:
        with _paths(py):
            import theano
            del theano.CALL_LOG[:]
            theano.BUDGET_S, theano._FIRST_CALL[0] = budget_s, None
            with contextlib.redirect_stdout(buf if quiet else sys.stdout):
                try:
                    g = runpy.run_path(os.path.join(REF, "FNN_wnzh.py"), run_name="__main__")
                except theano.BudgetExceeded:
                    g = {"__stopped__": True}
                finally:
                    theano.BUDGET_S = None
            calls = list(theano.CALL_LOG)
    finally:
        sys.argv = old_argv
    g["__stdout__"] = buf.getvalue()
    return g, calls


def run_snn_dae(workdir, weights, quiet=True, budget_s=None, script="SNN_DAE.py"):
    """Executes the converted SNN_DAE.py (or script="SNN_RBM.py") top to bottom.  workdir/data/{train,test}.fm.txt must exist
    (batch_size is 1000 and x_dim 133465 in the scripts, SNN_DAE.py:19,41); `weights` = (ww0, bb0, ww1, bb1, ww2, bb2) is written to
    the pickle the script loads instead of pre-training (dropda_2997_.p, SNN_DAE.py:83-86; rbm_2997_.p, SNN_RBM.py:84-86).  Returns
    (globals, calls) like run_fnn_wnzh."""
    import pickle
    assert available(), "oracle/_ref is not built (needs /root/reference)"
    py = os.path.join(workdir, "python")
    os.makedirs(py, exist_ok=True)
    os.makedirs(os.path.join(workdir, "log"), exist_ok=True)
    with open(os.path.join(py, {"SNN_DAE.py": "dropda_2997_.p", "SNN_RBM.py": "rbm_2997_.p"}[script]), "wb") as f:
        pickle.dump(tuple(weights), f)
    _fresh("theano", "theano.tensor", "theano.tensor.shared_randomstreams", "theano.sparse", "dl_utils", "data_fm",
           "sampling_based_denosing_autoencoder", "sampling_based_gaussian_binary_rbm_sparse")
    linecache.clearcache()
    old_argv = sys.argv
    sys.argv = [script]
    buf = io.StringIO()
    try:
        with _paths(py):
            import theano
            del theano.CALL_LOG[:]
            theano.BUDGET_S, theano._FIRST_CALL[0] = budget_s, None
            with contextlib.redirect_stdout(buf if quiet else sys.stdout):
                try:
                    g = runpy.run_path(os.path.join(REF, script), run_name="__main__")
                except theano.BudgetExceeded:
                    g = {"__stopped__": True}
                finally:
                    theano.BUDGET_S = None
            calls = list(theano.CALL_LOG)
    finally:
        sys.argv = old_argv
    g["__stdout__"] = buf.getvalue()
    return g, calls


_TF_NAMES = ("tf_util", "FM", "LR", "FNN", "FNN_IP_L3", "FNN_IP_L5", "FNN_IP_L7", "FNN_OP_L3")


def tf_modules(dtype=None):
    """dict name -> module (plus 'tf' = the shim).  dtype: torch.float32 for TF's own precision, default float64."""
    assert available(), "oracle/_ref is not built (needs /root/reference)"
    with _paths():
        import tensorflow as tf
        if dtype is not None:
            tf.SHIM_DTYPE = dtype
        mods = {"tf": tf}
        with contextlib.redirect_stdout(io.StringIO()):
            for n in _TF_NAMES:
                mods[n] = importlib.import_module(n)
    return mods


def load_baseline(workdir, quiet=True):
    """Imports the converted baseline.py with cwd = <workdir>/python (it appends its header to ../log/<tag> on import)."""
    assert available()
    py = os.path.join(workdir, "python")
    os.makedirs(py, exist_ok=True)
    os.makedirs(os.path.join(workdir, "log"), exist_ok=True)
    os.makedirs(os.path.join(workdir, "model"), exist_ok=True)
    _fresh("baseline")
    with _paths(py):
        with contextlib.redirect_stdout(io.StringIO() if quiet else sys.stdout):
            return importlib.import_module("baseline")


def load_ipinyou():
    assert available()
    _fresh("ipinyou")
    with _paths():
        with contextlib.redirect_stdout(io.StringIO()):
            return importlib.import_module("ipinyou")


@contextlib.contextmanager
def in_reference_cwd(workdir):
    """cwd = <workdir>/python while reference functions that use relative paths run."""
    with _paths(os.path.join(workdir, "python")):
        yield
