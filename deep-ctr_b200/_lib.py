"""ctypes binding of libdctr.so (include/dctr.h).  There is NO fallback: if the CUDA library is missing
or a call fails, this raises."""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DCTR_LIB") or os.path.join(_HERE, "libdctr.so")      # DCTR_LIB: an experimental build of the same ABI
CSRC = os.path.join(_HERE, "csrc")

OK, EINVAL, EWORKSPACE, ECUDA = 0, -1, -2, -3
ACT = {"none": 0, None: 0, "tanh": 1, "relu": 2, "sigmoid": 3}
CROSS = {"none": 0, None: 0, "ip": 1, "op": 2}
SCATTER_SUM, SCATTER_SEQ_DECAY = 0, 1
GEMM_NN, GEMM_NT, GEMM_TN = 0, 1, 2
EPI_NONE, EPI_FWD, EPI_DGRAD = 0, 1, 2

_p, _i, _l, _f, _z, _d = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t, ctypes.c_double

# name -> (restype, argtypes); must list every symbol include/dctr.h declares (tests/test_abi.py checks)
SIGNATURES = {
    "dctr_version": (_i, []),
    "dctr_last_error": (ctypes.c_char_p, []),
    "dctr_launch_count": (ctypes.c_longlong, []),
    "dctr_gather_concat_fwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _i, _i, _p, _i, _p, _i, _p]),
    "dctr_arrange_fields": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p]),
    "dctr_arrange_update_list": (_i, [_p, _p, _i, _i, _i, _l, _i, _p, _p, _p, _p, _p]),
    "dctr_set_id_status": (_i, [_p]),
    "dctr_gather_rows": (_i, [_p, _p, _l, _i, _l, _p, _p]),
    "dctr_ipc_alloc": (_i, [_z, _p, _p]),
    "dctr_ipc_open": (_i, [_p, _p]),
    "dctr_ipc_close": (_i, [_p]),
    "dctr_ipc_free": (_i, [_p]),
    "dctr_gather_rows_push": (_i, [_p, _p, _l, _i, _l, _i, _i, _p, _p]),
    "dctr_slot_grads_push": (_i, [_p, _p, _p, _p, _l, _i, _i, _i, _l, _i, _i, _p, _p]),
    "dctr_cross_bwd_push": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _i, _p, _i, _i, _p, _p, _l, _i, _p]),
    "dctr_shard_pack": (_i, [_p, _p, _p, _l, _i, _l, _p, _p, _p, _p, _p]),
    "dctr_shard_pack_push": (_i, [_p, _p, _p, _l, _i, _l, _p, _p, _p, _p, _i, _p, _p]),
    "dctr_host_pack_fields": (_i, [_p, _i, _p, _p, _p, _l, _i, _i, _p, _p, _i]),
    "dctr_cross_bwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _i, _p, _i, _p, _i, _p]),
    "dctr_cross_bwd_to": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _i, _p, _i, _p, _i, _p, _p]),
    "dctr_fm_fwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _p, _p, _f, _p, _p, _p, _p, _p, _i, _p]),
    "dctr_sort_rows_workspace": (_z, [_l]),
    "dctr_sort_rows": (_i, [_p, _l, _l, _p, _p, _p, _z, _p]),
    "dctr_sort_pairs": (_i, [_p, _p, _l, _l, _p, _p, _p, _z, _p]),
    "dctr_scatter_workspace": (_z, [_l, _i]),
    "dctr_scatter_sgd": (_i, [_i, _p, _p, _l, _i, _i, _p, _l, _i, _p, _f, _d, _p, _l, _i, _p, _z, _p]),
    "dctr_scatter_sgd_masked": (_i, [_i, _p, _p, _l, _i, _i, _p, _l, _i, _i, _f, _d, _p, _l, _i, _p, _z, _p]),
    "dctr_scatter_sgd_fm": (_i, [_p, _p, _l, _i, _i, _p, _p, _i, _p, _f, _d, _p, _l, _i, _p, _z, _p]),
    "dctr_scatter_grad_fm": (_i, [_p, _p, _l, _i, _i, _p, _p, _i, _p, _p, _l, _i, _p, _p, _z, _p]),
    "dctr_decay_sweep": (_i, [_p, _l, _i, _f, _p, _l, _p]),
    "dctr_gemm": (_i, [_i, _i, _i, _i, _p, _i, _p, _i, _p, _i, _i, _i, _p, _p, _i, _p, _i, _i, _p]),
    "dctr_splitk_reduce": (_i, [_p, _i, _i, _i, _i, _p, _i, _p]),
    "dctr_splitk_reduce_sgd": (_i, [_p, _i, _i, _i, _i, _p, _i, _f, _f, _p]),
    "dctr_gemm_tc": (_i, [_i, _i, _i, _i, _p, _i, _p, _i, _p, _i, _i, _i, _p, _p, _i, _p, _i, _i, _p]),
    "dctr_reduce_workspace": (_z, [_i, _i]),
    "dctr_out_xent_fwd": (_i, [_p, _i, _i, _i, _p, _p, _p, _f, _p, _p, _p, _p, _f, _p, _p, _z, _p]),
    "dctr_out_bwd_fused": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _i, _p, _p, _p, _z, _p]),
    "dctr_out_bwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _i, _p]),
    "dctr_col_wsum_workspace": (_z, [_i, _i]),
    "dctr_col_wsum": (_i, [_p, _i, _i, _i, _p, _p, _p, _z, _p]),
    "dctr_col_wsum_sgd": (_i, [_p, _i, _i, _i, _p, _p, _f, _f, _p, _z, _p]),
    "dctr_col_wsum_static_max_h": (_i, []),
    "dctr_col_wsum_static": (_i, [_p, _i, _i, _i, _p, _p, _p]),
    "dctr_vec_sum": (_i, [_p, _i, _f, _p, _p]),
    "dctr_set_launch_mode": (_i, [_i]),
    "dctr_eval_metrics_workspace": (_z, [_l]),
    "dctr_eval_metrics": (_i, [_p, _p, _l, _p, _p, _z, _p]),
    "dctr_out_unit_train": (_i, [_p, _i, _i, _i, _p, _p, _p, _f, _i, _p, _i, _p, _p, _p, _p, _f, _p, _p, _i, _p, _p, _p, _z, _p]),
    "dctr_dropout_masks": (_i, [_p, _l, _f, ctypes.c_uint, _p, _p, _l, _p]),
    "dctr_l2_loss_add": (_i, [_p, _l, _p, _i, _f, _p, _p, _z, _p]),
    "dctr_mlp2_transpose": (_i, [_i, _i, _i, _p, _i, _p, _i, _p, _i, _p, _i, _p]),
    "dctr_mlp2_supported": (_i, [_i, _i, _i]),
    "dctr_mlp2_workspace": (_z, [_i, _i, _i, _i]),
    "dctr_mlp2_train": (_i, [_i, _i, _i, _i, _p, _i, _p, _i, _i, _p, _i, _p, _i, _p, _i, _p, _i, _p, _i, _p, _i, _p, _p, _p, _p, _p, _f, _f,
                             _p, _i, _p, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _i, _f, _f, _p, _z, _p]),
    "dctr_mlp2tc_supported": (_i, [_i, _i, _i]),
    "dctr_mlp2tc_workspace": (_z, [_i, _i]),
    "dctr_mlp2tc_packed_bytes": (_z, [_i, _i, _i]),
    "dctr_mlp2tc_pack": (_i, [_i, _i, _i, _p, _i, _p, _i, _p, _z, _p]),
    "dctr_mlp2tc_train": (_i, [_i, _i, _i, _i, _p, _i, _p, _i, _i, _p, _p, _i, _p, _i, _p, _i, _p, _i, _p, _p, _p, _f, _f,
                               _p, _i, _p, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _i, _f, _f, _p, _z, _p]),
    "dctr_mlp2_wgrad": (_i, [_i, _i, _i, _i, _p, _i, _p, _i, _p, _p, _i, _p, _i, _p, _p, _i, _p, _p, _p, _p, _p, _i, _f, _f, _p, _z, _p]),
    "dctr_dense_sgd": (_i, [_p, _p, _l, _f, _f, _p]),
    "dctr_adam_tick": (_i, [_p, _f, _f, _f, _p]),
    "dctr_dense_adam": (_i, [_p, _p, _p, _p, _l, _p, _f, _f, _f, _f, _i, _p]),
    "dctr_dense_ftrl": (_i, [_p, _p, _p, _p, _l, _f, _f, _f, _f, _f, _i, _p]),
    "dctr_bag_fwd": (_i, [_p, _p, _p, _l, _i, _i, _i, _i, _p, _i, _p, _i, _p]),
    "dctr_text_index_workspace": (_z, [_l]),
    "dctr_text_index_lines": (_i, [_p, _l, _p, _l, _p, _p, _z, _p]),
    "dctr_text_parse_sparse": (_i, [_p, _l, _p, _p, _l, _i, _i, _i, _p, _p, _p, _p, _p]),
    "dctr_text_parse_criteo": (_i, [_p, _l, _p, _p, _l, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
}

_lib = None


def build(force=False, verbose=True):
    """Compile libdctr.so in-tree for sm_100a (nvcc cross-compiles without a GPU).  Prints which sources were compiled and which
    objects were reused (the prebuilt objects travel with the snapshot to the GPU box)."""
    if force:
        subprocess.run(["make", "-C", CSRC, "clean"], check=True, capture_output=True)
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libdctr.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        compiled = sorted(set(ln.split(" -c ")[1].split()[0] for ln in r.stdout.splitlines() if " -c " in ln))
        srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
        linked = any(" -shared " in ln for ln in r.stdout.splitlines())
        print("deepctr_b200.build: %d of %d CUDA sources compiled for sm_100a (%s), %d objects reused, libdctr.so %s"
              % (len(compiled), len(srcs), ", ".join(compiled) or "none", len(srcs) - len(compiled), "linked" if linked else "up to date"))
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libdctr.so is missing (%s): run `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`make -C deep-ctr_b200/csrc`.  There is no CPU fallback." % LIB_PATH)
        h = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        if h.dctr_version() != 2:
            raise RuntimeError("libdctr.so version mismatch")
        _lib = h
    return _lib


PROFILE = None   # bench.py sets this to a list; every call then appends (name, args, start_event, end_event)
NVTX = os.environ.get("DCTR_NVTX", "0") == "1"      # one NVTX range per C-ABI call (K0..K7 show up by name in nsys / ncu --nvtx)


def call(name, *args):
    """Invoke an int-returning entry point; raise RuntimeError with dctr_last_error() on failure."""
    if PROFILE is not None and name not in ("dctr_set_launch_mode", "dctr_host_pack_fields", "dctr_set_id_status"):      # host-only entry points
        import torch
        st = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        rc = getattr(lib(), name)(*args)
        e1.record(st)
        PROFILE.append((name, args, e0, e1))
    elif NVTX:
        import torch
        torch.cuda.nvtx.range_push(name)
        try:
            rc = getattr(lib(), name)(*args)
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        rc = getattr(lib(), name)(*args)
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (name, rc, lib().dctr_last_error().decode()))


def launch_count():
    return int(lib().dctr_launch_count())


def ptr(t):
    """device pointer of a torch tensor (or None)."""
    return None if t is None else t.data_ptr()
