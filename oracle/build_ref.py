"""Builds oracle/_ref/ -- the reference's OWN Python, made runnable here -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

/root/reference/python/*.py is Python 2 on Theano / TensorFlow 0.x; neither engine nor a python2 interpreter is in this
image.  The hot path's index logic, host loops and graph-BUILDING code are plain Python, so this script reads the
sources where they lie, applies mechanical Py2 -> Py3 fixes and writes the result to oracle/_ref/python/ (git-ignored:
reference sources are never committed; the directory travels to the GPU box like a built .so).  The converted files run
on the two small engine shims under oracle/shims/ (`theano`, `tensorflow`: lazy graphs evaluated by torch, autograd for
T.grad / minimize).  What runs is therefore the reference's own statements: feat_layer_one_index /
feats_to_layer_one_array / get_fxy / the triple update loop (FNN_wnzh.py:87-96, 248-253, 298-306), build_inds and
init_var_map (tf_util.py:7-12, 41-82), the forward() graph builders of FM / LR / FNN / FNN_IP_L3/L5/L7 / FNN_OP_L3,
get_fxy / get_batch_xy / early_stop / watch_train (baseline.py:190-281), collect / stat / load_ipinyou_data
(ipinyou.py:11-65), and the SNN fine-tuning scripts SNN_DAE.py / SNN_RBM.py (get_fi_h1_y SNN_DAE.py:234-260, the graph :113-163,
the ww0 / bb0 update loop :282-288).

Fixes applied (every one is asserted to match, so a changed reference fails loudly):
  * `print x, y` statements -> print(x, y)            (token-level, any file)
  * `import cPickle as pickle` -> `import pickle`
  * the PATCHES table below: Py2 integer division used as a loop bound, list-returning range(), text files opened 'rb'.
Nothing else is touched; no arithmetic, index or control-flow statement is edited.

Usage:  python oracle/build_ref.py            (no-op with a message when /root/reference is absent)
"""
import hashlib
import io
import json
import os
import sys
import tokenize

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("DCTR_REFERENCE", "/root/reference/python")
OUT = os.path.join(HERE, "_ref", "python")

FILES = ["FNN_wnzh.py", "dl_utils.py", "data_fm.py", "tf_util.py", "FM.py", "LR.py", "FNN.py", "FNN_IP_L3.py", "FNN_IP_L5.py",
         "FNN_IP_L7.py", "FNN_OP_L3.py", "baseline.py", "ipinyou.py",
         # the SNN fine-tuning script (embedding-bag first layer); its pre-training module is only imported
         # (get_da_weights / get_rbm_weights are not called when the weight pickle of SNN_DAE.py:84-86 / SNN_RBM.py:84-86 exists)
         "SNN_DAE.py", "sampling_based_denosing_autoencoder.py", "SNN_RBM.py", "sampling_based_gaussian_binary_rbm_sparse.py"]

# (file, old, new, why) -- `old` must occur exactly `count` times (default 1)
PATCHES = [
    ("FNN_wnzh.py", "n_batch = train_size / batch_size", "n_batch = train_size // batch_size",
     "Py2 int / int is floor division; the value bounds range()"),
    ("baseline.py", "_f = range(13)", "_f = list(range(13))", "Py2 range() returns a list (.extend follows)"),
    ("baseline.py", "data_set = open(train_path, 'rb')", "data_set = open(train_path, 'r')", "Py2 'rb' yields str lines"),
    ("baseline.py", "data_set = open(test_path, 'rb')", "data_set = open(test_path, 'r')", "Py2 'rb' yields str lines"),
    ("baseline.py", "range(labels.shape[0] / batch_size)", "range(labels.shape[0] // batch_size)", "Py2 floor division"),
    ("baseline.py", "range(eval_buf_size / eval_size)", "range(eval_buf_size // eval_size)", "Py2 floor division"),
    ("baseline.py", "range(labels.shape[0] / test_batch_size)", "range(labels.shape[0] // test_batch_size)", "Py2 floor division"),
    ("ipinyou.py", "inds = range(len(y))", "inds = list(range(len(y)))", "Py2 range() returns a list (shuffled in place)"),
    ("ipinyou.py", "open(train_path, 'rb')", "open(train_path, 'r')", "Py2 'rb' yields str lines"),
    ("ipinyou.py", "open(test_path, 'rb')", "open(test_path, 'r')", "Py2 'rb' yields str lines"),
    ("ipinyou.py", "range(len(batch_labels) / batch_size)", "range(len(batch_labels) // batch_size)", "Py2 floor division"),
    ("data_fm.py", None, None, None),
    ("SNN_DAE.py", "n_batch=train_size/batch_size", "n_batch=train_size//batch_size", "Py2 floor division; the value bounds range()"),
    ("SNN_DAE.py", "start_t = time.clock()", "start_t = time.perf_counter()", "time.clock() is gone in Python 3.8+ (an unused timer)"),
    ("SNN_RBM.py", "n_batch = train_size / batch_size", "n_batch = train_size // batch_size", "Py2 floor division; the value bounds range()"),
    ("SNN_RBM.py", "start_t = time.clock()", "start_t = time.perf_counter()", "time.clock() is gone in Python 3.8+ (an unused timer)"),
]


def convert_prints(src):
    """Rewrite Python-2 print STATEMENTS as calls.  Token-level: a NAME `print` that starts a statement and is not followed
    by `(`-only call syntax gets its argument list wrapped up to the end of the logical line."""
    toks = list(tokenize.generate_tokens(io.StringIO(src).readline))
    lines = src.splitlines(keepends=True)
    offs = [0]
    for ln in lines:
        offs.append(offs[-1] + len(ln))

    def pos(rc):
        return offs[rc[0] - 1] + rc[1]

    edits = []
    start_of_stmt = True
    i = 0
    while i < len(toks):
        t = toks[i]
        if t.type == tokenize.NAME and t.string == "print" and start_of_stmt:
            j = i + 1
            depth = 0
            while not (toks[j].type == tokenize.NEWLINE and depth == 0) and toks[j].type != tokenize.ENDMARKER:
                if toks[j].type == tokenize.OP and toks[j].string in "([{":
                    depth += 1
                elif toks[j].type == tokenize.OP and toks[j].string in ")]}":
                    depth -= 1
                j += 1
            last = toks[j - 1]
            if last.type == tokenize.COMMENT:
                last = toks[j - 2]
            a, b = pos(t.end), pos(last.end)
            body = src[a:b].strip()
            end = ""
            if body.endswith(","):
                body, end = body[:-1].rstrip(), ", end=' '"
            edits.append((pos(t.start), b, "print(%s%s)" % (body, end)))
            i = j
            start_of_stmt = True
            continue
        if t.type in (tokenize.NEWLINE, tokenize.INDENT, tokenize.DEDENT, tokenize.NL) or (t.type == tokenize.OP and t.string in ":;"):
            start_of_stmt = True
        elif t.type != tokenize.COMMENT:
            start_of_stmt = False
        i += 1
    for a, b, new in reversed(edits):
        src = src[:a] + new + src[b:]
    return src, len(edits)


def convert(name, src):
    notes = []
    src, n = convert_prints(src)
    if n:
        notes.append("%d print statements" % n)
    if "import cPickle as pickle" in src:
        src = src.replace("import cPickle as pickle", "import pickle")
        notes.append("cPickle -> pickle")
    for f, old, new, why in PATCHES:
        if f != name or old is None:
            continue
        cnt = src.count(old)
        if cnt != 1:
            raise RuntimeError("%s: expected exactly one %r, found %d -- the reference changed" % (name, old, cnt))
        src = src.replace(old, new)
        notes.append("%s -> %s (%s)" % (old, new, why))
    compile(src, name, "exec")          # must be valid Python 3
    return src, notes


def build(verbose=True):
    """Returns the output directory, or None when the reference tree is absent (GPU box: the prebuilt files travel)."""
    if not os.path.isdir(SRC):
        if verbose:
            print("build_ref: %s absent; %s" % (SRC, "using the prebuilt oracle/_ref" if os.path.isdir(OUT) else "oracle/_ref unavailable"))
        return OUT if os.path.isdir(OUT) else None
    os.makedirs(OUT, exist_ok=True)
    manifest = {"source": SRC, "files": {}}
    for name in FILES:
        with open(os.path.join(SRC, name)) as f:
            raw = f.read()
        out, notes = convert(name, raw)
        with open(os.path.join(OUT, name), "w") as f:
            f.write(out)
        manifest["files"][name] = {"sha256_source": hashlib.sha256(raw.encode()).hexdigest(), "fixes": notes}
    with open(os.path.join(HERE, "_ref", "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    if verbose:
        print("build_ref: %d reference files converted into %s" % (len(FILES), OUT))
    return OUT


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
