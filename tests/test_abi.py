"""CPU-side checks of the drop-in boundary: libdctr.so builds, loads and exports every symbol that
include/dctr.h declares, and argument validation fails loudly without touching a GPU."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "dctr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dctr_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(built_lib):
    from deepctr_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 20
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(raw, s), "libdctr.so does not export %s" % s
        assert s in _lib.SIGNATURES, "no ctypes signature for %s" % s
    assert set(_lib.SIGNATURES) == set(syms)


def test_version_and_error_slot(built_lib):
    from deepctr_b200 import _lib
    assert built_lib.dctr_version() == 2
    # bad arguments are rejected before any CUDA call: ld not a multiple of 4
    rc = built_lib.dctr_gather_concat_fwd(1, None, 1, 10, 3, 1, 2, 3, 0, 0, None, 0, 1, 16, None)
    assert rc == _lib.EINVAL
    assert b"gather_concat_fwd" in built_lib.dctr_last_error()
    try:
        _lib.call("dctr_gemm", 7, 1, 1, 1, 16, 4, 16, 4, 16, 4, 0, 0, None, None, 0, None, 0, 1, None)
    except RuntimeError as e:
        assert "bad layout" in str(e)
    else:
        raise AssertionError("expected RuntimeError")


def test_workspace_queries(built_lib):
    assert built_lib.dctr_sort_rows_workspace(65536) >= 2 * 65536 * 4
    assert built_lib.dctr_scatter_workspace(65536, 12) > 0
    assert built_lib.dctr_col_wsum_workspace(4096, 300) >= 300 * 4


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from deepctr_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    try:
        _lib.lib()
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("expected RuntimeError")


def test_host_pack_fields(built_lib):
    """the one HOST entry point: offset add + int32 cast + numeric | categorical concatenation (FNN.py:41-43, 65-69)"""
    import numpy as np
    from deepctr_b200 import _lib
    rng = np.random.default_rng(3)
    for N, nn, nc, dt in [(1, 0, 3, np.int64), (10001, 13, 26, np.int64), (50000, 2, 5, np.int32), (7, 3, 0, np.int64)]:
        c = rng.integers(0, 1000, (N, nc)).astype(dt)
        off = (np.arange(nc) * 1000 + nn).astype(np.int64)
        v = rng.random((N, nn)).astype(np.float32)
        cw = rng.random((N, nc)).astype(np.float32)
        ids = np.full((N, nn + nc), -1, np.int32)
        w = np.full((N, nn + nc), -1, np.float32)
        for wts_given in (True, False):
            _lib.call("dctr_host_pack_fields", c.ctypes.data if nc else None, c.dtype.itemsize, off.ctypes.data if nc else None,
                      v.ctypes.data if nn else None, cw.ctypes.data if (wts_given and nc) else None, N, nn, nc,
                      ids.ctypes.data, w.ctypes.data, 4)
            assert np.array_equal(ids, np.concatenate([np.broadcast_to(np.arange(nn), (N, nn)), c.astype(np.int64) + off], 1))
            assert np.array_equal(w, np.concatenate([v, cw if wts_given else np.ones_like(cw)], 1))
    assert built_lib.dctr_host_pack_fields(None, 2, None, None, None, 1, 0, 1, None, None, 1) == _lib.EINVAL


def test_every_device_entry_point_is_a_torch_op(built_lib):
    """north_star: "a thin C-ABI exposed as PyTorch custom ops" -- one torch.ops.deepctr.<name> per device entry point of
    include/dctr.h (host helpers, size queries and per-thread settings have no tensors to dispatch on), each declared with a
    schema that marks what it writes, CUDA-only."""
    import torch
    from deepctr_b200 import ops
    host_side = {"dctr_version", "dctr_last_error", "dctr_launch_count", "dctr_set_launch_mode", "dctr_set_id_status",
                 "dctr_host_pack_fields", "dctr_mlp2_supported", "dctr_mlp2tc_supported", "dctr_col_wsum_static_max_h",
                 "dctr_ipc_alloc", "dctr_ipc_open", "dctr_ipc_close", "dctr_ipc_free"}
    merged = {"dctr_sort_rows": "sort_pairs", "dctr_arrange_fields": "arrange_update_list", "dctr_scatter_sgd_masked": "scatter_sgd",
              "dctr_out_bwd": "out_bwd_fused"}      # entry points that are special cases of another one's op
    for sym in declared_symbols():
        if sym in host_side or sym.endswith("_workspace") or sym.endswith("_bytes"):
            continue
        name = merged.get(sym, sym[len("dctr_"):])
        assert name in ops.OPS, "no torch op for %s" % sym
        schema = str(getattr(torch.ops.deepctr, name).default._schema)
        assert "!" in schema, "%s declares no mutated argument: %s" % (name, schema)
    # no CPU fallback: the dispatcher has no CPU kernel to run
    try:
        torch.ops.deepctr.vec_sum(torch.zeros(4), 4, 1.0, torch.zeros(1))
    except NotImplementedError:
        pass
    else:
        raise AssertionError("a CPU tensor must not dispatch")
