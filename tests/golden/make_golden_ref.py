"""Generates tests/golden/ref_golden.pkl by EXECUTING THE REFERENCE (oracle/ref_run.py: /root/reference/python/*.py after the
mechanical Py2->3 fixes of oracle/build_ref.py, on the theano / tensorflow shims of oracle/shims/).  Unlike
oracle_golden.npz these vectors are outputs of the reference's own statements; the GPU parity tests
(tests/test_ref_golden_gpu.py) compare the CUDA path with them on a box that has no /root/reference.

  wnzh   the whole FNN_wnzh.py run on the S1 demo set (with a duplicate-field, a missing-field and a repeated-feature line):
         per train() call the feature lists, labels, dropout draws, gx and the weights it returned; per epoch the predictions
         of the train and test files; the final feat_weights table
  tf     FM, LR, FNN, FNN_IP_L3/L5/L7, FNN_OP_L3: 3 optimiser steps (sgd / adam / ftrl, dropout masks logged), eval graph, dump

Run:  python tests/golden/make_golden_ref.py     (needs /root/reference; every case is also asserted against the oracle)
"""
import os
import pathlib
import pickle
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import ref_cases as C  # noqa: E402
import ref_run  # noqa: E402
import test_oracle_vs_ref as T  # noqa: E402


def wnzh():
    wd = tempfile.mkdtemp()
    case = C.write_s1_demo(wd, dup_field=True)
    g, calls = ref_run.run_fnn_wnzh(wd)
    train = T.read_fm_lines(os.path.join(wd, "data", "train.fm.txt"))
    test = T.read_fm_lines(os.path.join(wd, "data", "test.fm.txt"))
    X = case["X"]
    rec = dict(table=case["table"], feat_field=case["feat_field"], w_0=case["w_0"], train=train, test=test,
               hyper={k: g[k] for k in ("batch_size", "lr", "dropout", "hidden1", "hidden2", "lambda1", "lambda_fm", "acti_type")},
               steps=[], epochs=[], final_table=np.array([g["feat_weights"][f] for f in range(X)]))
    tc = [c for c in calls if c["fn"] == "function1"]
    for i, c in enumerate(calls):
        if c["fn"] == "function1":
            st = dict(r1=c["random"]["binomial1"].astype(np.float32), r2=c["random"]["binomial2"].astype(np.float32))
            k = len(rec["steps"])
            if k < 6:                                      # the first two epochs step by step, the rest by their effect
                st.update(gx=c["outputs"][0].astype(np.float32), b1=c["outputs"][4], w3=c["outputs"][3])
            if k in (3, 5):                                # weights as train() returned them = after 3 / 5 updates
                st.update(w1=c["outputs"][1].astype(np.float32), w2=c["outputs"][2].astype(np.float32))
            rec["steps"].append(st)
    rec["min_err"], rec["min_err_epoch"] = float(g["min_err"]), int(g["min_err_epoch"])
    with open(os.path.join(wd, "python", "mlp3fm_test_2997.p"), "rb") as f:        # get_pred(test_file, best_w1, ...), FNN_wnzh.py:346-347
        rec["best_test_pred"] = np.asarray(pickle.load(f), dtype=np.float32)
    preds = [c["outputs"][0] for c in calls if c["fn"] == "function2"]
    rec["epochs"] = [dict(train_p=preds[2 * e].astype(np.float32), test_p=preds[2 * e + 1].astype(np.float32)) for e in range(len(preds) // 2 - 1)]
    return rec


def tf_cases():
    tfm = ref_run.tf_modules()
    out = {}
    plan = [("FM", "sgd", "sum"), ("FM", "adam", "mean"), ("LR", "sgd", "sum"), ("LR", "ftrl", "sum")]
    for name, pt, red in plan:
        rec = {}
        T.run_fm_case(tfm, pathlib.Path(tempfile.mkdtemp()), name, pt, red, record=rec)
        out["%s/%s/%s" % (name, pt, red)] = rec
    plan = [("FNN", "sgd", 0.5, "sum", "tanh"), ("FNN", "adam", 0.5, "sum", "relu"), ("FNN", "ftrl", 1.0, "mean", "sigmoid"),
            ("FNN_IP_L3", "sgd", 0.5, "sum", "tanh"), ("FNN_IP_L5", "sgd", 1.0, "mean", "relu"), ("FNN_IP_L7", "adam", 0.5, "sum", "tanh"),
            ("FNN_OP_L3", "sgd", 0.5, "sum", "tanh")]
    for name, pt, keep, red, act in plan:
        rec = {}
        T.run_fnn_case(tfm, pathlib.Path(tempfile.mkdtemp()), name, pt, keep, red, act, record=rec)
        rec.update(reduce=red, act=act)
        out["%s/%s/%g/%s/%s" % (name, pt, keep, red, act)] = rec
    return out


if __name__ == "__main__":
    assert ref_run.available(), "needs /root/reference"
    gold = dict(wnzh=wnzh(), tf=tf_cases())
    path = os.path.join(HERE, "ref_golden.pkl")
    with open(path, "wb") as f:
        pickle.dump(gold, f, protocol=4)
    print("wrote %s (%d bytes)" % (path, os.path.getsize(path)))
