"""Data-parallel training over one process per GPU (torch.distributed, NCCL over NVLink / NVSwitch).

The path shards by SAMPLES: every rank runs gather + MLP forward/backward on its N local
samples.  Two exchanges make the update identical on every rank and equal to a single-GPU step on the
concatenated (rank-major) global batch:
  dense   one all-reduce(SUM) of the flat MLP gradient buffer;
  sparse  all-gather of (ids, wts, per-sample embedding gradient); every rank then sorts the GLOBAL
          (row, position) list and folds it -- so replicated tables stay bit-identical across ranks and the
          sequential-decay semantics (FNN_wnzh.py:300-306) see the global sample order.
The row-sharded table with id / embedding / gradient all-to-all (10^8 rows) is the next row of 8(e).
The collectives below run on CPU tensors with gloo too, which is how tests/test_dist_cpu.py covers them.
"""
import torch

from . import _lib as L


def all_gather_cat(dist, t, group=None):
    """[n, ...] on every rank -> [world*n, ...] in rank order (same on every rank)."""
    world = dist.get_world_size(group)
    out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t.contiguous(), group=group)
    return out


def exchange_sparse(dist, ids, wts, g, group=None):
    """The sparse exchange: returns the global (ids, wts, g), rank-major."""
    return all_gather_cat(dist, ids, group), (None if wts is None else all_gather_cat(dist, wts, group)), all_gather_cat(dist, g, group)


def allreduce_dense(dist, flat, group=None):
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


class DataParallelStep(object):
    """Replaces StepEngine.train_step / FMEngine.train_step with the data-parallel step."""

    def __init__(self, engine, dist, group=None):
        self.e, self.dist, self.group = engine, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        e = engine
        self.GM = e.M * self.world
        i32 = dict(dtype=torch.int32, device=e.dev)
        self.g_rows = torch.zeros(self.GM, **i32)
        self.g_pos = torch.zeros(self.GM, **i32)
        self.sort_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(self.GM), dtype=torch.uint8, device=e.dev)
        self.scat_ws = torch.zeros(L.lib().dctr_scatter_workspace(self.GM, e.ld), dtype=torch.uint8, device=e.dev)
        self.is_fm = not hasattr(e, "widths")

    def _sort_global(self, gids):
        e = self.e
        L.call("dctr_sort_rows", gids.data_ptr(), self.GM, e.X, self.g_rows.data_ptr(), self.g_pos.data_ptr(),
               self.sort_ws.data_ptr(), self.sort_ws.numel(), torch.cuda.current_stream(e.dev).cuda_stream)

    def train_step(self, lr, decay=1.0, gscale=1.0, train_bias=True, last_decay=1.0):
        e, dist = self.e, self.dist
        cur = torch.cuda.current_stream(e.dev)
        gids = all_gather_cat(dist, e.ids, self.group)
        gwts = None if e.wts_are_ones else all_gather_cat(dist, e.wts, self.group)
        e.side.wait_stream(cur)
        with torch.cuda.stream(e.side):
            self._sort_global(gids)
        e.forward(True, gscale)
        dz = e.backward()
        st = cur.cuda_stream
        N, F, d = e.N, e.F, e.d
        ldd = dz.stride(0)
        if train_bias:
            col = 0 if e.bias_first else F * d + e.crossw
            L.call("dctr_col_wsum", dz.data_ptr() + 4 * col, ldd, N, 1, None, L.ptr(e.bias(e.grads)), L.ptr(e.col_ws), e.col_ws.numel(), st)
        allreduce_dense(dist, e.grads, self.group)
        allreduce_dense(dist, e.loss, self.group)
        if e.cross != "none":
            wts = None if e.wts_are_ones else e.wts
            L.call("dctr_cross_bwd", L.ptr(e.ids), L.ptr(wts), L.ptr(e.table), e.X, e.ld, N, F, d, L.CROSS[e.cross],
                   L.ptr(dz), ldd, L.ptr(e.ge), e.ldg, st)
            gg = all_gather_cat(dist, e.ge, self.group)
            g, g_ld_n, g_ld_f, w = gg.data_ptr(), F * e.ldg, e.ldg, None
        else:
            gg = all_gather_cat(dist, dz, self.group)
            g, g_ld_n, g_ld_f, w = gg.data_ptr() + (4 if e.bias_first else 0), ldd, d, L.ptr(gwts)
        cur.wait_stream(e.side)
        L.call("dctr_scatter_sgd", e.scatter_mode, self.g_rows.data_ptr(), self.g_pos.data_ptr(), self.GM, F, d, g, g_ld_n, g_ld_f, w,
               float(lr), float(decay), L.ptr(e.table), e.X, e.ld, self.scat_ws.data_ptr(), self.scat_ws.numel(), st)
        e.dense_update(lr, last_decay)
        self._keep = (gids, gwts, gg)       # keep the gathered buffers alive until the stream has consumed them

    def train_step_fm(self, lr, lam=0.0, gscale=1.0, dense_l2=True):
        e, dist = self.e, self.dist
        cur = torch.cuda.current_stream(e.dev)
        gids = all_gather_cat(dist, e.ids, self.group)
        gwts = None if e.wts_are_ones else all_gather_cat(dist, e.wts, self.group)
        e.side.wait_stream(cur)
        with torch.cuda.stream(e.side):
            self._sort_global(gids)
        e.forward(True, gscale)
        st = cur.cuda_stream
        L.call("dctr_vec_sum", L.ptr(e.g), e.N, 1.0, L.ptr(e.grads), st)
        allreduce_dense(dist, e.grads, self.group)
        allreduce_dense(dist, e.loss, self.group)
        gg, gs = all_gather_cat(dist, e.g, self.group), all_gather_cat(dist, e.s, self.group)
        cur.wait_stream(e.side)
        decay = 1.0 - lr * lam
        L.call("dctr_scatter_sgd_fm", self.g_rows.data_ptr(), self.g_pos.data_ptr(), self.GM, e.F, e.d, gg.data_ptr(), gs.data_ptr(),
               e.lds, L.ptr(gwts), float(lr), float(decay), L.ptr(e.table), e.X, e.ld, self.scat_ws.data_ptr(), self.scat_ws.numel(), st)
        if lam != 0.0 and dense_l2:
            L.call("dctr_decay_sweep", L.ptr(e.table), e.X, e.ld, float(decay), self.g_rows.data_ptr(), self.GM, st)
        L.call("dctr_dense_sgd", L.ptr(e.params), L.ptr(e.grads), 1, float(lr), float(decay), st)
        self._keep = (gids, gwts, gg, gs)


def attach(model, dist, group=None):
    """Switch a model (models.FM / FNN* / theano_fnn.FNN_wnzh) to data-parallel steps.  The reference's
    batch-size-dependent constants see the GLOBAL batch: 'mean' reduce and the decay 1-2*lambda*lr/b_size."""
    dp = DataParallelStep(model.engine, dist, group)
    model.dp = dp
    model.use_graph = False           # NCCL collectives are issued eagerly between the kernel groups
    model.global_N = model.N * dp.world
    if dp.is_fm:
        model.engine.train_step = dp.train_step_fm
    else:
        model.engine.train_step = dp.train_step
    return dp
