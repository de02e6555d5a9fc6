"""PyTorch custom ops over the C-ABI of libdctr.so: `torch.ops.deepctr.*` (north_star: "a thin C-ABI exposed as PyTorch
custom ops").  One op per device entry point of include/dctr.h, same name without the `dctr_` prefix.

Every op
  * is declared with torch.library (schema with `Tensor(a!)` on everything it writes: tables, parameters, outputs, workspaces),
  * has a CUDA implementation only -- there is no CPU fallback; on a CPU tensor the dispatcher raises NotImplementedError,
  * checks dtype, device, inner-stride-1 layout and 16-byte alignment where the kernels use 128-bit accesses, takes the
    leading dimensions from the tensors' strides (column-offset views such as dz[:, 1:] are passed as they are),
  * launches on torch.cuda.current_stream() of the tensors' device and raises RuntimeError with dctr_last_error() on failure.
The reference has no operator interface of its own (it builds Theano / TF graphs); the object protocol these ops sit under
is baseline.py:306-351 (models.py), the reference statement each op replaces is cited in include/dctr.h.
"""
import torch

from . import _lib as L

lib = torch.library.Library("deepctr", "DEF")
_F32, _I32, _U8, _I64, _F64 = torch.float32, torch.int32, torch.uint8, torch.int64, torch.float64
OPS = []          # names of every registered op (tests/test_abi.py checks them against include/dctr.h)


def _ptr(t, dtype, what, align=4):
    """device pointer of `t` after the boundary checks (None passes through)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise NotImplementedError("deepctr ops run on CUDA tensors only (%s is on %s): there is no CPU fallback" % (what, t.device))
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (what, dtype, t.dtype))
    if t.numel() > 1 and t.stride(-1) != 1:
        raise ValueError("%s: the innermost dimension must be contiguous (strides %r)" % (what, tuple(t.stride())))
    p = t.data_ptr()
    if p % align:
        raise ValueError("%s must be %d-byte aligned" % (what, align))
    return p


def _ld(t, what):
    """leading dimension (row stride in elements) of a 2-D view; one-row tensors broadcast with 0"""
    if t is None:
        return 0
    if t.dim() != 2:
        raise ValueError("%s must be 2-D, got shape %r" % (what, tuple(t.shape)))
    return 0 if t.shape[0] == 1 else t.stride(0)


def _st(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _op(name, schema):
    """decorator: declare deepctr::<name> with `schema` and register the function as its CUDA implementation"""
    def deco(fn):
        lib.define(name + schema)
        lib.impl(name, fn, "CUDA")
        OPS.append(name)
        return fn
    return deco


# ---- K0 / K1: arrange, gather + concat (+ IP / OP), row gather, cross-layer backward, embedding bag ---------------------------
@_op("arrange_update_list", "(Tensor feats, Tensor feat_field, int F, int X, int pad_row, Tensor(a!) ids, Tensor(b!) wts, "
     "Tensor(c!)? upd_rows, Tensor(d!)? upd_pos) -> ()")
def arrange_update_list(feats, feat_field, F, X, pad_row, ids, wts, upd_rows, upd_pos):
    N, J = feats.shape
    L.call("dctr_arrange_update_list", _ptr(feats, _I32, "feats"), _ptr(feat_field, _I32, "feat_field"), N, J, F, X, pad_row,
           _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(upd_rows, _I32, "upd_rows"), _ptr(upd_pos, _I32, "upd_pos"), _st(feats))


@_op("gather_concat_fwd", "(Tensor ids, Tensor? wts, Tensor table, int d, int cross, bool bias_first, Tensor bias, int act0, "
     "Tensor(a!) out) -> ()")
def gather_concat_fwd(ids, wts, table, d, cross, bias_first, bias, act0, out):
    N, F = ids.shape
    L.call("dctr_gather_concat_fwd", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0],
           table.stride(0), N, F, d, cross, int(bias_first), _ptr(bias, _F32, "bias"), act0, _ptr(out, _F32, "out", 16), out.stride(0), _st(ids))


@_op("gather_rows", "(Tensor ids, Tensor table, Tensor(a!) out) -> ()")
def gather_rows(ids, table, out):
    L.call("dctr_gather_rows", _ptr(ids, _I32, "ids"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0), ids.numel(),
           _ptr(out, _F32, "out", 16), _st(ids))


@_op("shard_pack", "(Tensor ids, Tensor perm, Tensor owner_sorted, int world, int C, Tensor(a!) send_ids, Tensor(b!) src_of_dest, "
     "Tensor(c!) inv, Tensor(d!) status) -> ()")
def shard_pack(ids, perm, owner_sorted, world, C, send_ids, src_of_dest, inv, status):
    if send_ids.numel() < world * C or src_of_dest.numel() < world * C or inv.numel() < ids.numel() or status.numel() < 2:
        raise ValueError("shard_pack: send_ids / src_of_dest need world * C entries, inv one per slot, status two words")
    L.call("dctr_shard_pack", _ptr(ids, _I32, "ids"), _ptr(perm, _I32, "perm"), _ptr(owner_sorted, _I32, "owner_sorted"), ids.numel(), world, C,
           _ptr(send_ids, _I32, "send_ids"), _ptr(src_of_dest, _I32, "src_of_dest"), _ptr(inv, _I32, "inv"), _ptr(status, _I32, "status"), _st(ids))


@_op("shard_pack_push", "(Tensor ids, Tensor perm, Tensor owner_sorted, int world, int C, Tensor(a!) send_ids, Tensor(b!) src_of_dest, "
     "Tensor(c!) inv, Tensor(d!) status, int rank, Tensor peers, Tensor(e!) own) -> ()")
def shard_pack_push(ids, perm, owner_sorted, world, C, send_ids, src_of_dest, inv, status, rank, peers, own):
    if send_ids.numel() < world * C or src_of_dest.numel() < world * C or inv.numel() < ids.numel() or status.numel() < 2:
        raise ValueError("shard_pack_push: send_ids / src_of_dest need world * C entries, inv one per slot, status two words")
    if peers.dtype != _I64 or peers.numel() < world or own.numel() < world * C:
        raise ValueError("shard_pack_push: peers = one int64 pointer per rank, each to world * C int32 entries")
    L.call("dctr_shard_pack_push", _ptr(ids, _I32, "ids"), _ptr(perm, _I32, "perm"), _ptr(owner_sorted, _I32, "owner_sorted"), ids.numel(), world, C,
           _ptr(send_ids, _I32, "send_ids"), _ptr(src_of_dest, _I32, "src_of_dest"), _ptr(inv, _I32, "inv"), _ptr(status, _I32, "status"),
           rank, _ptr(peers, _I64, "peers", 8), _st(ids))


@_op("cross_bwd", "(Tensor ids, Tensor? wts, Tensor table, int d, int cross, Tensor dz, Tensor(a!) ge) -> ()")
def cross_bwd(ids, wts, table, d, cross, dz, ge):
    N, F = ids.shape
    L.call("dctr_cross_bwd", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0),
           N, F, d, cross, _ptr(dz, _F32, "dz"), dz.stride(0), _ptr(ge, _F32, "ge"), ge.stride(1), _st(ids))


@_op("cross_bwd_to", "(Tensor ids, Tensor? wts, Tensor table, int d, int cross, Tensor dz, Tensor dest, Tensor(a!) ge) -> ()")
def cross_bwd_to(ids, wts, table, d, cross, dz, dest, ge):
    """ge: [rows, ldg]; slot (n, f) is written to row dest[n * F + f]"""
    N, F = ids.shape
    if dest.numel() < N * F or ge.dim() != 2:
        raise ValueError("cross_bwd_to: dest needs one entry per slot and ge must be [rows, ldg]")
    L.call("dctr_cross_bwd_to", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0),
           N, F, d, cross, _ptr(dz, _F32, "dz"), dz.stride(0), _ptr(ge, _F32, "ge"), ge.stride(0), _ptr(dest, _I32, "dest"), _st(ids))


@_op("gather_rows_push", "(Tensor ids, Tensor table, int C, int world, int rank, Tensor peers, Tensor(a!) own) -> ()")
def gather_rows_push(ids, table, C, world, rank, peers, own):
    """peers: int64 [world] device tensor of row-buffer base pointers (peers[rank] == own.data_ptr()); `own` only declares the write"""
    if ids.numel() < world * C or peers.numel() < world or peers.dtype != _I64:
        raise ValueError("gather_rows_push: ids needs world * C entries, peers one int64 pointer per rank")
    L.call("dctr_gather_rows_push", _ptr(ids, _I32, "ids"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0), C, world, rank,
           _ptr(peers, _I64, "peers", 8), _st(ids))


@_op("slot_grads_push", "(Tensor send_ids, Tensor src_of_dest, Tensor? wts, Tensor dz, int F, int d, int ldg, int C, int world, int rank, "
     "Tensor peers, Tensor(a!) own) -> ()")
def slot_grads_push(send_ids, src_of_dest, wts, dz, F, d, ldg, C, world, rank, peers, own):
    if send_ids.numel() < world * C or src_of_dest.numel() < world * C or peers.dtype != _I64 or peers.numel() < world:
        raise ValueError("slot_grads_push: send_ids / src_of_dest need world * C entries, peers one int64 pointer per rank")
    L.call("dctr_slot_grads_push", _ptr(send_ids, _I32, "send_ids"), _ptr(src_of_dest, _I32, "src_of_dest"), _ptr(wts, _F32, "wts"),
           _ptr(dz, _F32, "dz"), dz.stride(0), F, d, ldg, C, world, rank, _ptr(peers, _I64, "peers", 8), _st(dz))


@_op("cross_bwd_push", "(Tensor ids, Tensor? wts, Tensor table, int d, int cross, Tensor dz, Tensor dest, int ldg, int C, int rank, "
     "Tensor peers, Tensor(a!) own) -> ()")
def cross_bwd_push(ids, wts, table, d, cross, dz, dest, ldg, C, rank, peers, own):
    N, F = ids.shape
    if dest.numel() < N * F or peers.dtype != _I64:
        raise ValueError("cross_bwd_push: dest needs one entry per slot, peers one int64 pointer per rank")
    L.call("dctr_cross_bwd_push", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0),
           N, F, d, cross, _ptr(dz, _F32, "dz"), dz.stride(0), ldg, _ptr(dest, _I32, "dest"), _ptr(peers, _I64, "peers", 8), C, rank, _st(ids))


@_op("bag_fwd", "(Tensor ids, Tensor? wts, Tensor table, int H, Tensor? b0, int act, Tensor(a!) out) -> ()")
def bag_fwd(ids, wts, table, H, b0, act, out):
    N, F = ids.shape
    L.call("dctr_bag_fwd", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0),
           N, F, H, _ptr(b0, _F32, "b0"), act, _ptr(out, _F32, "out", 16), out.stride(0), _st(ids))


@_op("fm_fwd", "(Tensor ids, Tensor? wts, Tensor table, int d, Tensor bias, Tensor? y, float gscale, Tensor(a!) logit, "
     "Tensor(b!) pred, Tensor(c!) loss_n, Tensor(d!) g, Tensor(e!)? s) -> ()")
def fm_fwd(ids, wts, table, d, bias, y, gscale, logit, pred, loss_n, g, s):
    N, F = ids.shape
    L.call("dctr_fm_fwd", _ptr(ids, _I32, "ids"), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0),
           N, F, d, _ptr(bias, _F32, "bias"), _ptr(y, _F32, "y"), gscale, _ptr(logit, _F32, "logit"), _ptr(pred, _F32, "pred"),
           _ptr(loss_n, _F32, "loss_n"), _ptr(g, _F32, "g"), _ptr(s, _F32, "s", 16), 0 if s is None else s.stride(0), _st(ids))


# ---- K6: sort, folds, decay sweep ------------------------------------------------------------------------------------------------
@_op("sort_pairs", "(Tensor ids, Tensor? payload, int M, int X, Tensor(a!) sorted_rows, Tensor(b!) sorted_pos, Tensor(c!) workspace) -> ()")
def sort_pairs(ids, payload, M, X, sorted_rows, sorted_pos, workspace):
    L.call("dctr_sort_pairs", _ptr(ids, _I32, "ids"), _ptr(payload, _I32, "payload"), M, X, _ptr(sorted_rows, _I32, "sorted_rows", 16),
           _ptr(sorted_pos, _I32, "sorted_pos", 16), _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(ids))


def _g_lds(g, F, g_ld_f):
    """gradient source g: 2-D [N, >= F * g_ld_f] (per-sample rows, field f at column f * g_ld_f) or 3-D [N, F, ldg]"""
    if g.dim() == 3:
        return g.stride(0), g.stride(1)
    return g.stride(0), g_ld_f


@_op("scatter_sgd", "(int mode, Tensor sorted_rows, Tensor sorted_pos, int M, int F, int d, Tensor g, int g_ld_f, Tensor? wts, "
     "int skip_row, float lr, float decay, Tensor(a!) table, Tensor(b!) workspace) -> ()")
def scatter_sgd(mode, sorted_rows, sorted_pos, M, F, d, g, g_ld_f, wts, skip_row, lr, decay, table, workspace):
    """skip_row >= 0 with wts = None is dctr_scatter_sgd_masked; otherwise dctr_scatter_sgd"""
    ld_n, ld_f = _g_lds(g, F, g_ld_f)
    common = (_ptr(sorted_rows, _I32, "sorted_rows", 16), _ptr(sorted_pos, _I32, "sorted_pos", 16), M, F, d, _ptr(g, _F32, "g"), ld_n, ld_f)
    tail = (lr, decay, _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0), _ptr(workspace, _U8, "workspace", 16),
            workspace.numel(), _st(table))
    if skip_row >= 0:
        if wts is not None:
            raise ValueError("scatter_sgd: skip_row and wts are exclusive")
        L.call("dctr_scatter_sgd_masked", mode, *common, skip_row, *tail)
    else:
        L.call("dctr_scatter_sgd", mode, *common, _ptr(wts, _F32, "wts"), *tail)


@_op("scatter_sgd_fm", "(Tensor sorted_rows, Tensor sorted_pos, int M, int F, int d, Tensor g, Tensor s, Tensor? wts, float lr, "
     "float decay, Tensor(a!) table, Tensor(b!) workspace) -> ()")
def scatter_sgd_fm(sorted_rows, sorted_pos, M, F, d, g, s, wts, lr, decay, table, workspace):
    L.call("dctr_scatter_sgd_fm", _ptr(sorted_rows, _I32, "sorted_rows", 16), _ptr(sorted_pos, _I32, "sorted_pos", 16), M, F, d,
           _ptr(g, _F32, "g"), _ptr(s, _F32, "s", 16), s.stride(0), _ptr(wts, _F32, "wts"), lr, decay, _ptr(table, _F32, "table", 16),
           table.shape[0], table.stride(0), _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(table))


@_op("scatter_grad_fm", "(Tensor sorted_rows, Tensor sorted_pos, int M, int F, int d, Tensor g, Tensor s, Tensor? wts, Tensor table, "
     "Tensor(a!) grad_out, Tensor(b!) workspace) -> ()")
def scatter_grad_fm(sorted_rows, sorted_pos, M, F, d, g, s, wts, table, grad_out, workspace):
    L.call("dctr_scatter_grad_fm", _ptr(sorted_rows, _I32, "sorted_rows", 16), _ptr(sorted_pos, _I32, "sorted_pos", 16), M, F, d,
           _ptr(g, _F32, "g"), _ptr(s, _F32, "s", 16), s.stride(0), _ptr(wts, _F32, "wts"), _ptr(table, _F32, "table", 16), table.shape[0],
           table.stride(0), _ptr(grad_out, _F32, "grad_out", 16), _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(table))


@_op("decay_sweep", "(Tensor(a!) table, float decay, Tensor sorted_rows, int M) -> ()")
def decay_sweep(table, decay, sorted_rows, M):
    L.call("dctr_decay_sweep", _ptr(table, _F32, "table", 16), table.shape[0], table.stride(0), decay, _ptr(sorted_rows, _I32, "sorted_rows"),
           M, _st(table))


# ---- K4 / K5: GEMMs with fused epilogues, output unit, reductions ------------------------------------------------------------------
def _gemm(fn, layout, M, N, K, A, B, C, epilogue, act, bias, aux, mask, splits):
    L.call(fn, layout, M, N, K, _ptr(A, _F32, "A", 16), A.stride(0), _ptr(B, _F32, "B", 16), B.stride(0), _ptr(C, _F32, "C", 16),
           C.stride(0) if C.dim() == 2 else C.stride(1), epilogue, act, _ptr(bias, _F32, "bias"), _ptr(aux, _F32, "aux"),
           0 if aux is None else aux.stride(0), _ptr(mask, _F32, "mask"), _ld(mask, "mask"), splits, _st(A))


_GEMM_SCHEMA = ("(int layout, int M, int N, int K, Tensor A, Tensor B, Tensor(a!) C, int epilogue, int act, Tensor? bias, Tensor? aux, "
                "Tensor? mask, int splits) -> ()")


@_op("gemm", _GEMM_SCHEMA)
def gemm(layout, M, N, K, A, B, C, epilogue, act, bias, aux, mask, splits):
    _gemm("dctr_gemm", layout, M, N, K, A, B, C, epilogue, act, bias, aux, mask, splits)


@_op("gemm_tc", _GEMM_SCHEMA)
def gemm_tc(layout, M, N, K, A, B, C, epilogue, act, bias, aux, mask, splits):
    _gemm("dctr_gemm_tc", layout, M, N, K, A, B, C, epilogue, act, bias, aux, mask, splits)


@_op("splitk_reduce", "(Tensor partials, int splits, int M, int N, int ldc, Tensor(a!) out) -> ()")
def splitk_reduce(partials, splits, M, N, ldc, out):
    L.call("dctr_splitk_reduce", _ptr(partials, _F32, "partials"), splits, M, N, ldc, _ptr(out, _F32, "out"), out.stride(0), _st(out))


@_op("splitk_reduce_sgd", "(Tensor partials, int splits, int M, int N, int ldc, Tensor(a!) W, float lr, float decay) -> ()")
def splitk_reduce_sgd(partials, splits, M, N, ldc, W, lr, decay):
    L.call("dctr_splitk_reduce_sgd", _ptr(partials, _F32, "partials"), splits, M, N, ldc, _ptr(W, _F32, "W"), W.stride(0), lr, decay, _st(W))


@_op("out_xent_fwd", "(Tensor a, int H, Tensor w, Tensor b, Tensor? y, float gscale, Tensor(a!) logit, Tensor(b!) pred, Tensor(c!) loss_n, "
     "Tensor(d!) dlogit, float loss_scale, Tensor(e!)? loss_sum, Tensor(f!) workspace) -> ()")
def out_xent_fwd(a, H, w, b, y, gscale, logit, pred, loss_n, dlogit, loss_scale, loss_sum, workspace):
    L.call("dctr_out_xent_fwd", _ptr(a, _F32, "a"), a.stride(0), a.shape[0], H, _ptr(w, _F32, "w"), _ptr(b, _F32, "b"), _ptr(y, _F32, "y"), gscale,
           _ptr(logit, _F32, "logit"), _ptr(pred, _F32, "pred"), _ptr(loss_n, _F32, "loss_n"), _ptr(dlogit, _F32, "dlogit"), loss_scale,
           _ptr(loss_sum, _F32, "loss_sum"), _ptr(workspace, _U8, "workspace"), workspace.numel(), _st(a))


@_op("out_bwd_fused", "(Tensor dlogit, Tensor w, Tensor a, int H, int act, Tensor? mask, Tensor(a!) da, Tensor(b!) gw, Tensor(c!) gb, "
     "Tensor(d!) workspace) -> ()")
def out_bwd_fused(dlogit, w, a, H, act, mask, da, gw, gb, workspace):
    L.call("dctr_out_bwd_fused", _ptr(dlogit, _F32, "dlogit"), _ptr(w, _F32, "w"), _ptr(a, _F32, "a"), a.stride(0), a.shape[0], H, act,
           _ptr(mask, _F32, "mask"), _ld(mask, "mask"), _ptr(da, _F32, "da"), da.stride(0), _ptr(gw, _F32, "gw"), _ptr(gb, _F32, "gb"),
           _ptr(workspace, _U8, "workspace"), workspace.numel(), _st(a))


@_op("out_unit_train", "(Tensor a, int H, Tensor w, Tensor b, Tensor y, float gscale, int act, Tensor? mask, Tensor(a!) logit, "
     "Tensor(b!) pred, Tensor(c!) loss_n, Tensor(d!) dlogit, float loss_scale, Tensor(e!) loss_sum, Tensor(f!) da, Tensor(g!) gw, "
     "Tensor(h!) gb, Tensor(i!) workspace) -> ()")
def out_unit_train(a, H, w, b, y, gscale, act, mask, logit, pred, loss_n, dlogit, loss_scale, loss_sum, da, gw, gb, workspace):
    L.call("dctr_out_unit_train", _ptr(a, _F32, "a"), a.stride(0), a.shape[0], H, _ptr(w, _F32, "w"), _ptr(b, _F32, "b"), _ptr(y, _F32, "y"),
           gscale, act, _ptr(mask, _F32, "mask"), _ld(mask, "mask"), _ptr(logit, _F32, "logit"), _ptr(pred, _F32, "pred"),
           _ptr(loss_n, _F32, "loss_n"), _ptr(dlogit, _F32, "dlogit"), loss_scale, _ptr(loss_sum, _F32, "loss_sum"), _ptr(da, _F32, "da"),
           da.stride(0), _ptr(gw, _F32, "gw"), _ptr(gb, _F32, "gb"), _ptr(workspace, _U8, "workspace"), workspace.numel(), _st(a))


@_op("col_wsum", "(Tensor A, int H, Tensor? wv, Tensor(a!) out, Tensor(b!) workspace) -> ()")
def col_wsum(A, H, wv, out, workspace):
    L.call("dctr_col_wsum", _ptr(A, _F32, "A"), A.stride(0), A.shape[0], H, _ptr(wv, _F32, "wv"), _ptr(out, _F32, "out"),
           _ptr(workspace, _U8, "workspace"), workspace.numel(), _st(A))


@_op("col_wsum_sgd", "(Tensor A, int H, Tensor? wv, Tensor(a!) param, float lr, float decay, Tensor(b!) workspace) -> ()")
def col_wsum_sgd(A, H, wv, param, lr, decay, workspace):
    L.call("dctr_col_wsum_sgd", _ptr(A, _F32, "A"), A.stride(0), A.shape[0], H, _ptr(wv, _F32, "wv"), _ptr(param, _F32, "param"), lr, decay,
           _ptr(workspace, _U8, "workspace"), workspace.numel(), _st(A))


@_op("col_wsum_static", "(Tensor A, int H, Tensor? wv, Tensor(a!) out) -> ()")
def col_wsum_static(A, H, wv, out):
    """dctr_col_wsum on the library's built-in scratch (H <= 1024): nothing is allocated, calls on one device must not overlap."""
    L.call("dctr_col_wsum_static", _ptr(A, _F32, "A"), A.stride(0), A.shape[0], H, _ptr(wv, _F32, "wv"), _ptr(out, _F32, "out"), _st(A))


@_op("vec_sum", "(Tensor v, int N, float scale, Tensor(a!) out) -> ()")
def vec_sum(v, N, scale, out):
    L.call("dctr_vec_sum", _ptr(v, _F32, "v"), N, scale, _ptr(out, _F32, "out"), _st(v))


@_op("l2_loss_add", "(Tensor v, Tensor? extra, float scale, Tensor(a!) out, Tensor(b!) workspace) -> ()")
def l2_loss_add(v, extra, scale, out, workspace):
    L.call("dctr_l2_loss_add", _ptr(v, _F32, "v", 16), v.numel(), _ptr(extra, _F32, "extra"), 0 if extra is None else extra.numel(), scale,
           _ptr(out, _F32, "out"), _ptr(workspace, _U8, "workspace", 8), workspace.numel(), _st(v))


# ---- K7: dense optimizer sweeps ----------------------------------------------------------------------------------------------------
@_op("dense_sgd", "(Tensor(a!) param, Tensor grad, int n, float lr, float decay) -> ()")
def dense_sgd(param, grad, n, lr, decay):
    L.call("dctr_dense_sgd", _ptr(param, _F32, "param"), _ptr(grad, _F32, "grad"), n, lr, decay, _st(param))


@_op("adam_tick", "(Tensor(a!) state, float lr, float beta1, float beta2) -> ()")
def adam_tick(state, lr, beta1, beta2):
    L.call("dctr_adam_tick", _ptr(state, _F32, "state"), lr, beta1, beta2, _st(state))


@_op("dense_adam", "(Tensor(a!) param, Tensor(b!) grad, Tensor(c!) m, Tensor(d!) v, int n, Tensor state, float beta1, float beta2, float eps, "
     "float l2, bool zero_grad) -> ()")
def dense_adam(param, grad, m, v, n, state, beta1, beta2, eps, l2, zero_grad):
    L.call("dctr_dense_adam", _ptr(param, _F32, "param", 16), _ptr(grad, _F32, "grad", 16), _ptr(m, _F32, "m", 16), _ptr(v, _F32, "v", 16), n,
           _ptr(state, _F32, "state"), beta1, beta2, eps, l2, int(zero_grad), _st(param))


@_op("dense_ftrl", "(Tensor(a!) param, Tensor(b!) grad, Tensor(c!) accum, Tensor(d!) linear, int n, float lr, float lr_power, float l1, float l2, "
     "float grad_l2, bool zero_grad) -> ()")
def dense_ftrl(param, grad, accum, linear, n, lr, lr_power, l1, l2, grad_l2, zero_grad):
    L.call("dctr_dense_ftrl", _ptr(param, _F32, "param", 16), _ptr(grad, _F32, "grad", 16), _ptr(accum, _F32, "accum", 16),
           _ptr(linear, _F32, "linear", 16), n, lr, lr_power, l1, l2, grad_l2, int(zero_grad), _st(param))


@_op("dropout_masks", "(Tensor(a!) masks, float keep, int seed, Tensor(b!) state, Tensor(c!)? apply) -> ()")
def dropout_masks(masks, keep, seed, state, apply):
    L.call("dctr_dropout_masks", _ptr(masks, _F32, "masks", 16), masks.numel(), keep, seed, _ptr(state, _I32, "state"),
           _ptr(apply, _F32, "apply", 16), 0 if apply is None else apply.numel(), _st(masks))


# ---- fused two-hidden-layer step (csrc/mlp2.cu) ------------------------------------------------------------------------------------
@_op("mlp2_transpose", "(int D0, int H1, int H2, Tensor W1, Tensor W2, Tensor(a!) W1t, Tensor(b!) W2t) -> ()")
def mlp2_transpose(D0, H1, H2, W1, W2, W1t, W2t):
    L.call("dctr_mlp2_transpose", D0, H1, H2, _ptr(W1, _F32, "W1"), W1.stride(0), _ptr(W2, _F32, "W2"), W2.stride(0), _ptr(W1t, _F32, "W1t"),
           W1t.stride(0), _ptr(W2t, _F32, "W2t"), W2t.stride(0), _st(W1))


@_op("mlp2_train", "(int D0, int H1, int H2, Tensor(a!) x, Tensor? m0, int act0, Tensor W1, Tensor b1, int act1, Tensor? m1, Tensor W2, Tensor b2, "
     "int act2, Tensor? m2, Tensor W1t, Tensor W2t, Tensor(b!) w3, Tensor(c!) b3, Tensor y, float gscale, float loss_scale, Tensor(d!) a1, "
     "Tensor(e!) dz1, Tensor(f!) dz2, Tensor(g!) dz0, Tensor(h!) logit, Tensor(i!) pred, Tensor(j!) loss_n, Tensor(k!) dlogit, "
     "Tensor(l!) loss_sum, Tensor(m!) gw3, Tensor(n!) gb3, bool sgd, float lr, float decay_out, Tensor(o!) workspace) -> ()")
def mlp2_train(D0, H1, H2, x, m0, act0, W1, b1, act1, m1, W2, b2, act2, m2, W1t, W2t, w3, b3, y, gscale, loss_scale, a1, dz1, dz2, dz0,
               logit, pred, loss_n, dlogit, loss_sum, gw3, gb3, sgd, lr, decay_out, workspace):
    if dz1.stride(0) != dz2.stride(0):
        raise ValueError("mlp2_train: dz1 and dz2 must share their leading dimension")
    L.call("dctr_mlp2_train", x.shape[0], D0, H1, H2, _ptr(x, _F32, "x", 16), x.stride(0), _ptr(m0, _F32, "m0", 16), _ld(m0, "m0"), act0,
           _ptr(W1, _F32, "W1", 16), W1.stride(0), _ptr(b1, _F32, "b1"), act1, _ptr(m1, _F32, "m1"), _ld(m1, "m1"),
           _ptr(W2, _F32, "W2", 16), W2.stride(0), _ptr(b2, _F32, "b2"), act2, _ptr(m2, _F32, "m2"), _ld(m2, "m2"),
           _ptr(W1t, _F32, "W1t", 16), _ptr(W2t, _F32, "W2t", 16), _ptr(w3, _F32, "w3"), _ptr(b3, _F32, "b3"), _ptr(y, _F32, "y"), gscale, loss_scale,
           _ptr(a1, _F32, "a1", 16), a1.stride(0), _ptr(dz1, _F32, "dz1", 16), _ptr(dz2, _F32, "dz2", 16), dz1.stride(0), _ptr(dz0, _F32, "dz0", 16),
           _ptr(logit, _F32, "logit"), _ptr(pred, _F32, "pred"), _ptr(loss_n, _F32, "loss_n"), _ptr(dlogit, _F32, "dlogit"),
           _ptr(loss_sum, _F32, "loss_sum"), _ptr(gw3, _F32, "gw3"), _ptr(gb3, _F32, "gb3"), int(sgd), lr, decay_out,
           _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(x))


@_op("mlp2_wgrad", "(int D0, int H1, int H2, Tensor x, Tensor a1, Tensor dz1, Tensor dz2, Tensor(a!) W1, Tensor(b!) b1, Tensor(c!) W2, "
     "Tensor(d!) b2, Tensor(e!) gW1, Tensor(f!) gb1, Tensor(g!) gW2, Tensor(h!) gb2, bool sgd, float lr, float decay, Tensor(i!) workspace) -> ()")
def mlp2_wgrad(D0, H1, H2, x, a1, dz1, dz2, W1, b1, W2, b2, gW1, gb1, gW2, gb2, sgd, lr, decay, workspace):
    L.call("dctr_mlp2_wgrad", x.shape[0], D0, H1, H2, _ptr(x, _F32, "x", 16), x.stride(0), _ptr(a1, _F32, "a1", 16), a1.stride(0),
           _ptr(dz1, _F32, "dz1", 16), _ptr(dz2, _F32, "dz2", 16), dz1.stride(0), _ptr(W1, _F32, "W1"), W1.stride(0), _ptr(b1, _F32, "b1"),
           _ptr(W2, _F32, "W2"), W2.stride(0), _ptr(b2, _F32, "b2"), _ptr(gW1, _F32, "gW1"), _ptr(gb1, _F32, "gb1"), _ptr(gW2, _F32, "gW2"),
           _ptr(gb2, _F32, "gb2"), int(sgd), lr, decay, _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(x))


# ---- the fused two-hidden-layer step on the tensor cores (csrc/mlp2tc.cu) ------------------------------------------------------------
@_op("mlp2tc_pack", "(int D0, int H1, int H2, Tensor W1, Tensor W2, Tensor(a!) packed) -> ()")
def mlp2tc_pack(D0, H1, H2, W1, W2, packed):
    L.call("dctr_mlp2tc_pack", D0, H1, H2, _ptr(W1, _F32, "W1"), W1.stride(0), _ptr(W2, _F32, "W2"), W2.stride(0),
           _ptr(packed, _U8, "packed", 128), packed.numel(), _st(W1))


@_op("mlp2tc_train", "(int D0, int H1, int H2, Tensor(a!) x, Tensor? m0, int act0, Tensor packed, Tensor b1, int act1, Tensor? m1, "
     "Tensor b2, int act2, Tensor? m2, Tensor(b!) w3, Tensor(c!) b3, Tensor y, float gscale, float loss_scale, "
     "Tensor(d!) a1, Tensor(e!) dz1, Tensor(f!) dz2, Tensor(g!) dz0, Tensor(h!) logit, Tensor(i!) pred, Tensor(j!) loss_n, Tensor(k!) dlogit, "
     "Tensor(l!) loss_sum, Tensor(m!) gw3, Tensor(n!) gb3, bool sgd, float lr, float decay_out, Tensor(o!) workspace) -> ()")
def mlp2tc_train(D0, H1, H2, x, m0, act0, packed, b1, act1, m1, b2, act2, m2, w3, b3, y, gscale, loss_scale, a1, dz1, dz2, dz0,
                 logit, pred, loss_n, dlogit, loss_sum, gw3, gb3, sgd, lr, decay_out, workspace):
    if dz1.stride(0) != dz2.stride(0):
        raise ValueError("mlp2tc_train: dz1 and dz2 must share their leading dimension")
    if packed.numel() < L.lib().dctr_mlp2tc_packed_bytes(D0, H1, H2):
        raise ValueError("mlp2tc_train: packed weight buffer is smaller than dctr_mlp2tc_packed_bytes")
    L.call("dctr_mlp2tc_train", x.shape[0], D0, H1, H2, _ptr(x, _F32, "x", 16), x.stride(0), _ptr(m0, _F32, "m0", 16), _ld(m0, "m0"), act0,
           _ptr(packed, _U8, "packed", 128), _ptr(b1, _F32, "b1"), act1, _ptr(m1, _F32, "m1", 16), _ld(m1, "m1"),
           _ptr(b2, _F32, "b2"), act2, _ptr(m2, _F32, "m2", 16), _ld(m2, "m2"),
           _ptr(w3, _F32, "w3"), _ptr(b3, _F32, "b3"), _ptr(y, _F32, "y"), gscale, loss_scale,
           _ptr(a1, _F32, "a1", 16), a1.stride(0), _ptr(dz1, _F32, "dz1", 16), _ptr(dz2, _F32, "dz2", 16), dz1.stride(0), _ptr(dz0, _F32, "dz0", 16),
           _ptr(logit, _F32, "logit"), _ptr(pred, _F32, "pred"), _ptr(loss_n, _F32, "loss_n"), _ptr(dlogit, _F32, "dlogit"),
           _ptr(loss_sum, _F32, "loss_sum"), _ptr(gw3, _F32, "gw3"), _ptr(gb3, _F32, "gb3"), int(sgd), lr, decay_out,
           _ptr(workspace, _U8, "workspace", 16), workspace.numel(), _st(x))


# ---- eval metrics and device-side text ingestion -----------------------------------------------------------------------------------
@_op("eval_metrics", "(Tensor pred, Tensor y, Tensor(a!) out, Tensor(b!) workspace) -> ()")
def eval_metrics(pred, y, out, workspace):
    L.call("dctr_eval_metrics", _ptr(pred, _F32, "pred"), _ptr(y, _F32, "y"), pred.numel(), _ptr(out, _F64, "out", 8),
           _ptr(workspace, _U8, "workspace", 8), workspace.numel(), _st(pred))


@_op("text_index_lines", "(Tensor text, int nbytes, Tensor(a!) line_start, int max_lines, Tensor(b!) n_lines, Tensor(c!) workspace) -> ()")
def text_index_lines(text, nbytes, line_start, max_lines, n_lines, workspace):
    L.call("dctr_text_index_lines", _ptr(text, _U8, "text", 16), nbytes, _ptr(line_start, _I64, "line_start", 8), max_lines,
           _ptr(n_lines, _I64, "n_lines", 8), _ptr(workspace, _U8, "workspace", 8), workspace.numel(), _st(text))


@_op("text_parse_sparse", "(Tensor text, int nbytes, Tensor line_start, Tensor n_lines, int max_lines, int skip, int J, int pad_id, "
     "Tensor(a!) feats, Tensor(b!)? wts, Tensor(c!) y, Tensor(d!) status) -> ()")
def text_parse_sparse(text, nbytes, line_start, n_lines, max_lines, skip, J, pad_id, feats, wts, y, status):
    L.call("dctr_text_parse_sparse", _ptr(text, _U8, "text", 16), nbytes, _ptr(line_start, _I64, "line_start", 8), _ptr(n_lines, _I64, "n_lines", 8),
           max_lines, skip, J, pad_id, _ptr(feats, _I32, "feats"), _ptr(wts, _F32, "wts"), _ptr(y, _F32, "y"), _ptr(status, _I32, "status"), _st(text))


@_op("text_parse_criteo", "(Tensor text, int nbytes, Tensor line_start, Tensor n_lines, int max_lines, int n_num, int n_cat, Tensor offsets, "
     "Tensor max_vals, Tensor(a!) ids, Tensor(b!) wts, Tensor(c!) y, Tensor(d!) status) -> ()")
def text_parse_criteo(text, nbytes, line_start, n_lines, max_lines, n_num, n_cat, offsets, max_vals, ids, wts, y, status):
    L.call("dctr_text_parse_criteo", _ptr(text, _U8, "text", 16), nbytes, _ptr(line_start, _I64, "line_start", 8), _ptr(n_lines, _I64, "n_lines", 8),
           max_lines, n_num, n_cat, _ptr(offsets, _I64, "offsets", 8), _ptr(max_vals, _F64, "max_vals", 8), _ptr(ids, _I32, "ids"),
           _ptr(wts, _F32, "wts"), _ptr(y, _F32, "y"), _ptr(status, _I32, "status"), _st(text))
