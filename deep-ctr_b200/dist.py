"""Data-parallel training over one process per GPU (torch.distributed, NCCL over NVLink / NVSwitch).

The path shards by SAMPLES: every rank runs gather + MLP forward/backward on its N local
samples.  Two exchanges make the update identical on every rank and equal to a single-GPU step on the
concatenated (rank-major) global batch:
  dense   one all-reduce(SUM) of the flat MLP gradient buffer;
  sparse  all-gather of (ids, wts, per-sample embedding gradient); every rank then sorts the GLOBAL
          (row, position) list and folds it -- so replicated tables stay bit-identical across ranks and the
          sequential-decay semantics (FNN_wnzh.py:300-306) see the global sample order.
For vocabularies too large to replicate (10^8 rows, BASELINE config 5) the table is ROW-SHARDED instead
(ShardedStep): row g lives on rank g % world as local row g // world, and a step exchanges ids, embedding rows and
gradient rows with three all-to-alls; the owner folds the received gradients in (source rank, sample, field) order,
which is the global sample order, so the result equals the single-GPU step on the concatenated batch.
The collectives below run on CPU tensors with gloo too, which is how tests/test_dist_cpu.py covers them.
"""
import os

import torch

from . import _lib as L
from . import ops as _ops      # noqa: F401  (registers torch.ops.deepctr)

O = torch.ops.deepctr


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
    """Replaces StepEngine.train_step / FMEngine.train_step with the data-parallel step.  Every exchanged buffer is
    allocated once, so the whole step -- kernels and NCCL collectives -- is captured into one CUDA graph by the model."""

    def __init__(self, engine, dist, group=None):
        self.e, self.dist, self.group = engine, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        e = engine
        W = self.world
        self.GM = e.M * W
        i32 = dict(dtype=torch.int32, device=e.dev)
        f32 = dict(dtype=torch.float32, device=e.dev)
        self.g_rows = torch.zeros(self.GM, **i32)
        self.g_pos = torch.zeros(self.GM, **i32)
        self.sort_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(self.GM), dtype=torch.uint8, device=e.dev)
        self.scat_ws = torch.zeros(L.lib().dctr_scatter_workspace(self.GM, e.ld), dtype=torch.uint8, device=e.dev)
        self.is_fm = not hasattr(e, "widths")
        self.gids = torch.zeros(W * e.N, e.F, **i32)
        self.gwts = torch.zeros(W * e.N, e.F, **f32)
        self.loss_slot = e.grads[e.n_params:e.n_params + 1]
        self.ar_stream = torch.cuda.Stream(device=e.dev)
        if self.is_fm:
            self.gg, self.gs = torch.zeros(W * e.N, **f32), torch.zeros(W * e.N, e.lds, **f32)
        elif e.cross != "none":
            if e.ge is None:
                e.ge = torch.zeros(e.N, e.F, e.ldg, **f32)
            self.gg = torch.zeros(W * e.N, e.F, e.ldg, **f32)
        else:
            self.gg = torch.zeros(W * e.N, (e.dz0 if e.nl > 1 else e.dbuf[0]).stride(0), **f32)

    def _gather(self, out, t):
        self.dist.all_gather_into_tensor(out, t.contiguous(), group=self.group)
        return out

    def _sort_global(self, gids):
        e = self.e
        O.sort_pairs(gids, None, self.GM, e.X, self.g_rows, self.g_pos, self.sort_ws)

    def _reduce_dense(self):
        """ONE all-reduce for the dense gradients and the loss (the loss rides in the tail of the gradient buffer)."""
        e = self.e
        self.loss_slot.copy_(e.loss)
        allreduce_dense(self.dist, e.grads, self.group)
        e.loss.copy_(self.loss_slot)

    def train_step(self, lr, decay=1.0, gscale=1.0, train_bias=True, last_decay=1.0):
        e = self.e
        L.call("dctr_set_launch_mode", 0)          # collectives sit between the kernels: plain stream order
        cur = torch.cuda.current_stream(e.dev)
        e.side.wait_stream(cur)
        with torch.cuda.stream(e.side):          # id / weight exchange and the global sort: off the forward's critical path
            gids = self._gather(self.gids, e.ids)
            # weights reach the fold only when they are not 0 / 1-with-the-0s-on-the-pad-row (engine.skip_row: masked fold)
            gwts = None if (e.wts_are_ones or e.skip_row is not None) else self._gather(self.gwts, e.wts)
            self._sort_global(gids)
        e.forward(True, gscale)
        dz = e.backward()
        N, F, d = e.N, e.F, e.d
        if train_bias:
            col = 0 if e.bias_first else F * d + e.crossw
            O.col_wsum(dz[:, col:col + 1], 1, None, e.bias(e.grads), e.col_ws)
        if e.cross != "none":
            wts = None if e.wts_are_ones else e.wts
            O.cross_bwd(e.ids.view(N, F), None if wts is None else wts.view(N, F), e.table, d, L.CROSS[e.cross], dz, e.ge)
            g, g_ld_f, w = self._gather(self.gg, e.ge), e.ldg, None
        else:
            g, g_ld_f, w = self._gather(self.gg, dz)[:, (1 if e.bias_first else 0):], d, gwts
        # the dense all-reduce (issued after the gradient all-gather: collectives keep their issue order) runs on its own stream next
        # to the fold, which only needs the gathered slot gradients and the sorted global list
        self.ar_stream.wait_stream(cur)
        with torch.cuda.stream(self.ar_stream):
            e.join_wgrad()
            self._reduce_dense()
        cur.wait_stream(e.side)
        e.table_update(self.g_rows, self.g_pos, self.GM, F, g, g_ld_f, w, lr, decay, self.scat_ws)
        cur.wait_stream(self.ar_stream)
        e.dense_update(lr, last_decay)

    def train_step_fm(self, lr, lam=0.0, gscale=1.0, dense_l2=True):
        e = self.e
        L.call("dctr_set_launch_mode", 0)          # collectives sit between the kernels: plain stream order
        cur = torch.cuda.current_stream(e.dev)
        e.side.wait_stream(cur)
        with torch.cuda.stream(e.side):          # id / weight exchange and the global sort: off the forward's critical path
            gids = self._gather(self.gids, e.ids)
            gwts = None if e.wts_are_ones else self._gather(self.gwts, e.wts)
            self._sort_global(gids)
        e.forward(True, gscale, lam if self.rank == 0 else 0.0)      # the loss is summed over the ranks: the L2 term enters once
        O.vec_sum(e.g, e.N, 1.0, e.grads)
        self._reduce_dense()
        e.bias_update(lr, lam)
        gg, gs = self._gather(self.gg, e.g), self._gather(self.gs, e.s)
        cur.wait_stream(e.side)
        e.apply_update(self.g_rows, self.g_pos, self.GM, gg, gs, gwts, lr, lam, dense_l2, self.scat_ws)


# ======================================================================================================
# row-sharded embedding table (third row)
# ======================================================================================================
def plan_exchange(flat_ids, world, stable_argsort):
    """Bucket the (sample, field) slots of one rank by owner.  flat_ids: [M] global row ids.
    Returns perm [M] (slots ordered by owner rank, original order kept inside a bucket -- `stable_argsort(keys)` must
    be stable), counts [world] (slots per owner) and send_rows [M] (owner-local row ids, in perm order)."""
    owner = flat_ids % world
    perm = stable_argsort(owner)
    counts = torch.bincount(owner.long(), minlength=world)
    send_rows = torch.div(flat_ids, world, rounding_mode="floor")[perm.long()]
    return perm, counts, send_rows


def all_to_all_rows(dist, send, send_counts, recv_counts, group=None):
    """Variable-split all-to-all of the leading dimension; counts are Python ints per rank."""
    out = torch.empty((int(sum(recv_counts)),) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(out, send.contiguous(), output_split_sizes=list(recv_counts), input_split_sizes=list(send_counts), group=group)
    return out


def exchange_counts(dist, counts, group=None):
    """counts[r] = slots this rank sends to rank r  ->  recv[r] = slots rank r sends to this rank."""
    recv = torch.empty_like(counts)
    dist.all_to_all_single(recv, counts, group=group)
    return recv


class ShardedStep(object):
    """Train / predict step over a row-sharded embedding table.  The engine's `table` holds only this rank's rows
    (global row g = local row g // world on rank g % world); `engine.ids` carries GLOBAL row ids.

      1  bucket the N*F slots by owner (stable: dctr_sort_rows on the owner key), all-to-all the counts and the ids
      2  owners gather the requested rows (dctr_gather_rows), all-to-all them back
      3  gather_concat / MLP forward + backward on the local samples, reading the exchanged row buffer
      4  dense gradients + loss: all-reduce;  per-slot embedding gradients (dctr_cross_bwd): all-to-all to the owners
      5  owners sort the received local rows and fold (dctr_scatter_sgd) -- received order is (source rank, sample,
         field) = global sample order, so SUM and SEQ_DECAY both match the single-GPU step on the concatenated batch
    """

    def __init__(self, engine, dist, X_global, group=None):
        self.e, self.dist, self.group = engine, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.X_global = int(X_global)
        e = engine
        assert not e.bias_first, "row-sharded mode covers the TF-form models (bias column last)"
        assert e.X >= (self.X_global + self.world - 1) // self.world, "engine table must hold ceil(X / world) local rows"
        i32 = dict(dtype=torch.int32, device=e.dev)
        self.owner_sorted = torch.zeros(e.M, **i32)
        self.perm = torch.zeros(e.M, **i32)
        self.inv = torch.zeros(e.N, e.F, **i32)
        self.arange = torch.arange(e.M, **i32)
        self.plan_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(e.M), dtype=torch.uint8, device=e.dev)
        self.ge = e.ge if e.ge is not None else torch.zeros(e.N, e.F, e.ldg, dtype=torch.float32, device=e.dev)
        self.last = None
        # Fixed-capacity exchanges (DCTR_SHARD_FIXED=0 keeps the variable-split path): the first step runs with the split sizes read
        # on the host and measures the fullest (source, owner) bucket over all ranks; from the second step on every bucket has C =
        # slack x that entries, the three all-to-alls have equal splits, every buffer is static and the step replays as a CUDA graph.
        self.fixed = os.environ.get("DCTR_SHARD_FIXED", "1") != "0"
        self.slack = float(os.environ.get("DCTR_SHARD_SLACK", "1.15"))
        self.C = None
        self.overflow = torch.zeros(2, **i32)       # [slots that did not fit a bucket since the last check, fullest bucket seen]
        self._prof = [] if os.environ.get("DCTR_SHARD_PROFILE") else None
        self._train = False
        self._ar_pending = False
        self.ar_stream = torch.cuda.Stream(device=e.dev)

    def _mark(self, name):
        """DCTR_SHARD_PROFILE=1 (eager steps only): CUDA events between the phases of a step; summary() prints the averages"""
        if self._prof is None or torch.cuda.is_current_stream_capturing():
            return
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self._prof.append((name, ev))

    def summary(self):
        if not self._prof:
            return {}
        torch.cuda.synchronize()
        acc, cnt = {}, {}
        for (n0, e0), (n1, e1) in zip(self._prof[:-1], self._prof[1:]):
            if n1 == "start":
                continue
            acc[n1] = acc.get(n1, 0.0) + e0.elapsed_time(e1)
            cnt[n1] = cnt.get(n1, 0) + 1
        self._prof = []
        return {k: round(acc[k] / cnt[k], 4) for k in acc}

    def set_capacity(self, C):
        """allocate the static exchange buffers for buckets of C slots per (source, owner) pair"""
        e = self.e
        C = (int(C) + 255) // 256 * 256
        RC = self.world * C
        i32 = dict(dtype=torch.int32, device=e.dev)
        f32 = dict(dtype=torch.float32, device=e.dev)
        self.C, self.RC = C, RC
        self.send_ids, self.recv_ids, self.src_of_dest = torch.zeros(RC, **i32), torch.zeros(RC, **i32), torch.zeros(RC, **i32)
        self.gather_ids = torch.zeros(RC, **i32)
        self.p2p = self._setup_p2p(RC) if os.environ.get("DCTR_SHARD_P2P", "1") != "0" and self.world > 1 else False
        if not self.p2p:
            self.rows_buf, self.back = torch.zeros(RC, e.ld, **f32), torch.zeros(RC, e.ld, **f32)
            self.send_g, self.recv_g = torch.zeros(RC, e.ldg, **f32), torch.zeros(RC, e.ldg, **f32)
        self.srows, self.spos = torch.zeros(RC, **i32), torch.zeros(RC, **i32)
        self.sws = torch.zeros(L.lib().dctr_sort_rows_workspace(RC), dtype=torch.uint8, device=e.dev)
        self.fws = torch.zeros(L.lib().dctr_scatter_workspace(RC, e.ld), dtype=torch.uint8, device=e.dev)

    def _setup_p2p(self, RC):
        """The two row-sized exchanges as kernels that store into peer memory: `back` (rows the owners deliver) and `recv_g` (per-slot
        gradients the requesters deliver) live in cudaMalloc'ed buffers every rank of the node maps through CUDA IPC.  Returns False
        (and the step keeps the NCCL all-to-alls) if any rank fails to allocate or map."""
        import ctypes
        import sys
        e, dist = self.e, self.dist
        mine, ok = [], True
        try:
            for cols in (e.ld, e.ldg, 1):
                ptr, h = ctypes.c_void_p(), (ctypes.c_ubyte * 64)()
                L.call("dctr_ipc_alloc", RC * cols * 4, ctypes.byref(ptr), h)
                mine.append((ptr.value, bytes(h)))
        except Exception as ex:                                      # noqa: BLE001
            ok, err = False, str(ex)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (ok, [h for _, h in mine]), group=self.group)
        peers = [[0] * self.world for _ in range(3)]
        if all(o for o, _ in everyone):
            try:
                for r, (_, handles) in enumerate(everyone):
                    for b in range(3):
                        if r == self.rank:
                            peers[b][r] = mine[b][0]
                        else:
                            ptr = ctypes.c_void_p()
                            L.call("dctr_ipc_open", (ctypes.c_ubyte * 64).from_buffer_copy(handles[b]), ctypes.byref(ptr))
                            peers[b][r] = ptr.value
            except Exception as ex:                                  # noqa: BLE001
                ok, err = False, str(ex)
        else:
            ok, err = False, "a rank could not allocate"
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=e.dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if not int(flag.item()):
            if self.rank == 0:
                print("deepctr_b200.dist: peer-memory exchange unavailable (%s); using NCCL all-to-alls" % (err if not ok else "another rank failed"),
                      file=sys.stderr)
            for ptr, _ in mine:
                L.call("dctr_ipc_free", ptr)
            return False

        class _Dev(object):                                          # torch view of memory this module owns (never freed before exit)
            def __init__(self, ptr, shape, typestr="<f4"):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None}
        self.back = torch.as_tensor(_Dev(mine[0][0], (RC, e.ld)), device=e.dev)
        self.recv_g = torch.as_tensor(_Dev(mine[1][0], (RC, e.ldg)), device=e.dev)
        self.back_peers = torch.tensor(peers[0], dtype=torch.int64, device=e.dev)
        self.recvg_peers = torch.tensor(peers[1], dtype=torch.int64, device=e.dev)
        # the id buckets travel the same way (dctr_shard_pack_push) unless DCTR_SHARD_P2P_IDS=0 keeps the id all-to-all
        self.p2p_ids = os.environ.get("DCTR_SHARD_P2P_IDS", "1") != "0"
        if self.p2p_ids:
            # (not initialised here: a peer may already be storing its first buckets; every entry is written by its source each step)
            self.recv_ids = torch.as_tensor(_Dev(mine[2][0], (RC,), "<i4"), device=e.dev)
            self.ids_peers = torch.tensor(peers[2], dtype=torch.int64, device=e.dev)
        self.bar = torch.zeros(1, dtype=torch.float32, device=e.dev)
        return True

    def check_overflow(self):
        """host side, after a step: a bucket that overflowed dropped slots -- the step is invalid"""
        if self.C is None:
            return
        lost = int(self.overflow[0].item())
        if lost:
            self.overflow.zero_()
            C = self.C
            self.C = None                          # the next step re-measures
            raise RuntimeError("row-sharded exchange: %d slot(s) did not fit the buckets of %d entries per rank pair (the id distribution "
                               "changed since the calibration step); this step's update is incomplete -- raise DCTR_SHARD_SLACK" % (lost, C))

    def _st(self):
        return torch.cuda.current_stream(self.e.dev).cuda_stream

    def _stable_argsort(self, owner):
        e = self.e
        O.sort_pairs(owner.contiguous(), None, e.M, self.world, self.owner_sorted, self.perm, self.plan_ws)
        return self.perm

    def lookup(self):
        """steps 1-2: returns (inv ids [N,F], exchanged rows [M, ld]) and remembers the plan for the backward."""
        e, dist = self.e, self.dist
        if self.C is not None:
            return self._lookup_fixed()
        perm, counts, send_rows = plan_exchange(e.ids.view(-1), self.world, self._stable_argsort)
        recv_counts = exchange_counts(dist, counts, self.group)
        sc, rc = counts.tolist(), recv_counts.tolist()                    # host sync: the split sizes
        recv_rows = all_to_all_rows(dist, send_rows.to(torch.int32), sc, rc, self.group)
        R = recv_rows.shape[0]
        rows = torch.empty(max(R, 1), e.ld, dtype=torch.float32, device=e.dev)
        if R:
            O.gather_rows(recv_rows, e.table, rows)
        back = all_to_all_rows(dist, rows[:R], rc, sc, self.group)        # [M, ld], in perm order
        self.inv.view(-1)[perm.long()] = self.arange                      # slot -> its row in `back`
        self.last = (sc, rc, recv_rows, back)
        if self.fixed:                                                    # calibration: the fullest bucket over all ranks and pairs
            mx = torch.maximum(counts.max(), recv_counts.max()).reshape(1)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
            self._want_C = int(mx.item() * self.slack) + 64
        return self.inv, back

    def _lookup_fixed(self):
        """steps 1-2 with equal-sized buckets: no host read, static buffers"""
        e, dist = self.e, self.dist
        ids = e.ids.view(-1)
        self._mark("start")
        owner = torch.remainder(ids, self.world).to(torch.int32)
        O.sort_pairs(owner, None, e.M, self.world, self.owner_sorted, self.perm, self.plan_ws)
        self._mark("owner sort")
        if self.p2p and self.p2p_ids:      # the id buckets are stored into the owners' request lists; the all-reduce is the barrier
            O.shard_pack_push(ids, self.perm, self.owner_sorted, self.world, self.C, self.send_ids, self.src_of_dest, self.inv.view(-1),
                              self.overflow, self.rank, self.ids_peers, self.recv_ids)
            self._mark("pack")
            dist.all_reduce(self.bar, group=self.group)
        else:
            O.shard_pack(ids, self.perm, self.owner_sorted, self.world, self.C, self.send_ids, self.src_of_dest, self.inv.view(-1), self.overflow)
            self._mark("pack")
            dist.all_to_all_single(self.recv_ids, self.send_ids, group=self.group)
        self._mark("a2a ids")
        if self._train:      # the owner-side sort of the received rows only needs the ids: on the side stream, under the forward / backward
            main = torch.cuda.current_stream(e.dev)
            e.side.wait_stream(main)
            with torch.cuda.stream(e.side):
                # the unused entries (-1) become the sort's sentinel: behind every table row, never read or written by the fold
                L.call("dctr_set_id_status", None)
                O.sort_pairs(self.recv_ids, None, self.RC, e.X, self.srows, self.spos, self.sws)
                L.call("dctr_set_id_status", self.status_ptr)
        if self.p2p:
            # lookup + delivery in one kernel: every requested row is stored into the requester's buffer over NVLink (unused entries
            # are not transmitted); the small all-reduce is the barrier after which every rank's `back` is complete
            O.gather_rows_push(self.recv_ids, e.table, self.C, self.world, self.rank, self.back_peers, self.back)
            self._mark("owner gather")
            dist.all_reduce(self.bar, group=self.group)
            self._mark("a2a rows")
        else:
            torch.clamp(self.recv_ids, min=0, out=self.gather_ids)            # unused entries (-1) read row 0; nobody looks at them
            O.gather_rows(self.gather_ids, e.table, self.rows_buf)
            self._mark("owner gather")
            dist.all_to_all_single(self.back, self.rows_buf, group=self.group)
            self._mark("a2a rows")
        self.last = None
        return self.inv, self.back

    def forward(self, with_loss, gscale=1.0, train=False):
        """collective inference / forward: every rank must call it together (the lookups are all-to-alls)."""
        self._train = train
        inv, back = self.lookup()
        self.e.forward(with_loss, gscale, lookup=(inv, back, max(back.shape[0], 1)))

    def train_step(self, lr, decay=1.0, gscale=1.0, train_bias=True, last_decay=1.0):
        e, dist = self.e, self.dist
        L.call("dctr_set_launch_mode", 0)          # collectives sit between the kernels: plain stream order
        self.forward(True, gscale, train=True)
        if self.last is None:
            return self._train_tail_fixed(lr, decay, train_bias, last_decay)
        sc, rc, recv_rows, back = self.last
        dz = e.backward()
        N, F, d = e.N, e.F, e.d
        if train_bias:
            col = F * d + e.crossw
            O.col_wsum(dz[:, col:col + 1], 1, None, e.bias(e.grads), e.col_ws)
        e.join_wgrad()
        allreduce_dense(dist, e.grads, self.group)
        allreduce_dense(dist, e.loss, self.group)
        # per-slot embedding gradient (weights applied), against the exchanged pre-update rows
        wts = None if e.wts_are_ones else e.wts
        O.cross_bwd(self.inv.view(N, F), None if wts is None else wts.view(N, F), back, d, L.CROSS[e.cross], dz, self.ge)
        send_g = torch.empty(e.M, e.ldg, dtype=torch.float32, device=e.dev)
        O.gather_rows(self.perm, self.ge.view(e.M, e.ldg), send_g)
        recv_g = all_to_all_rows(dist, send_g, sc, rc, self.group)       # aligned with recv_rows
        R = recv_rows.shape[0]
        if R == 0 and e.dense_optimizer:      # dense Adam / FTRL moves every local row even when no gradient arrived
            e.table_update(None, None, 0, 1, recv_g, e.ldg, None, lr, decay, e.scat_ws)
        if R > 0:
            i32 = dict(dtype=torch.int32, device=e.dev)
            rows, pos = torch.empty(R, **i32), torch.empty(R, **i32)
            sws = torch.empty(L.lib().dctr_sort_rows_workspace(R), dtype=torch.uint8, device=e.dev)
            fws = torch.empty(L.lib().dctr_scatter_workspace(R, e.ld), dtype=torch.uint8, device=e.dev)
            O.sort_pairs(recv_rows, None, R, e.X, rows, pos, sws)
            # one "field" per received slot: position p indexes recv_g directly
            e.table_update(rows, pos, R, 1, recv_g, e.ldg, None, lr, decay, fws)
            self._keep = (rows, pos, sws, fws, recv_g, send_g, recv_rows, back)
        e.dense_update(lr, last_decay)
        if self.fixed and self.C is None:
            self.set_capacity(self._want_C)           # from the next step on: fixed-capacity exchanges

    def _train_tail_fixed(self, lr, decay, train_bias, last_decay):
        e, dist = self.e, self.dist
        self._mark("forward")
        dz = e.backward()                     # data gradients on this stream, weight gradients on the engine's second stream
        N, F, d = e.N, e.F, e.d
        if train_bias:
            col = F * d + e.crossw
            O.col_wsum(dz[:, col:col + 1], 1, None, e.bias(e.grads), e.col_ws)
        self._mark("backward")
        # the embedding side first: it only needs dz, and the weight-gradient GEMMs run next to it
        wts = None if e.wts_are_ones else e.wts
        # per-slot gradients written straight into the owner buckets (unused entries keep stale rows: the owner drops them)
        if self.p2p:      # ... of the OWNER's gradient buffer, over NVLink; the all-reduce is the barrier before the owners fold
            if e.cross == "none":     # no cross term: walk the buckets, every remote store streams contiguous rows
                O.slot_grads_push(self.send_ids, self.src_of_dest, None if wts is None else wts.view(-1), dz, F, d, e.ldg, self.C, self.world,
                                  self.rank, self.recvg_peers, self.recv_g)
            else:
                O.cross_bwd_push(self.inv.view(N, F), None if wts is None else wts.view(N, F), self.back, d, L.CROSS[e.cross], dz,
                                 self.inv.view(-1), e.ldg, self.C, self.rank, self.recvg_peers, self.recv_g)
            self._mark("slot grads")
            # the owner-side sort joins BEFORE the barrier: a peer that has passed it may start its next step and overwrite recv_ids
            torch.cuda.current_stream(e.dev).wait_stream(e.side)
            dist.all_reduce(self.bar, group=self.group)
            # the dense gradients' all-reduce only needs the weight gradients: issued AFTER the barrier (collectives keep their issue
            # order) on its own stream, it runs next to the owners' fold instead of behind it
            main = torch.cuda.current_stream(e.dev)
            self.ar_stream.wait_stream(main)
            with torch.cuda.stream(self.ar_stream):
                e.join_wgrad()
                allreduce_dense(dist, e.grads, self.group)
                allreduce_dense(dist, e.loss, self.group)
            self._ar_pending = True
        else:
            O.cross_bwd_to(self.inv.view(N, F), None if wts is None else wts.view(N, F), self.back, d, L.CROSS[e.cross], dz, self.inv.view(-1),
                           self.send_g)
            self._mark("slot grads")
            dist.all_to_all_single(self.recv_g, self.send_g, group=self.group)             # aligned with recv_ids
        self._mark("a2a grads")
        torch.cuda.current_stream(e.dev).wait_stream(e.side)                           # the sorted (row, position) list of the received slots
        e.table_update(self.srows, self.spos, self.RC, 1, self.recv_g, e.ldg, None, lr, decay, self.fws)
        self._mark("fold")
        if self._ar_pending:
            torch.cuda.current_stream(e.dev).wait_stream(self.ar_stream)
            self._ar_pending = False
        else:
            e.join_wgrad()
            allreduce_dense(dist, e.grads, self.group)
            allreduce_dense(dist, e.loss, self.group)
        self._mark("allreduce dense")
        e.dense_update(lr, last_decay)
        self._mark("dense update")


def attach_sharded(model, dist, X_global, group=None):
    """Switch a TF-form model built with `shard=(rank, world)` (its engine table holds the local rows only) to the
    row-sharded step.  'mean' reduce sees the GLOBAL batch; forward / predict become collective calls."""
    from .engine import BagEngine
    if isinstance(model.engine, BagEngine):
        raise NotImplementedError("the SNN embedding-bag step has no row-sharded form yet")
    sh = ShardedStep(model.engine, dist, X_global, group)
    model.sharded = sh
    sh.status_ptr = model._status_ptr
    # variable-split exchanges read their split sizes on the host every step; with fixed-capacity buckets (the default) the first
    # step calibrates and the later ones replay as a CUDA graph, collectives included (DCTR_SHARD_GRAPH=0: eager)
    model.use_graph = sh.fixed and os.environ.get("DCTR_SHARD_GRAPH", "1") != "0"
    model._graph, model._warm = None, 0
    check_ids = model._check_ids

    def check_ids_and_buckets():
        check_ids()
        sh.check_overflow()
    model._check_ids = check_ids_and_buckets
    model.global_N = model.N * sh.world
    model.engine.train_step = sh.train_step
    eng_forward = model.engine.forward

    def forward(with_loss, gscale=1.0, lookup=None):
        if lookup is not None:
            return eng_forward(with_loss, gscale, lookup)
        return sh.forward(with_loss, gscale)
    model.engine.forward = forward
    return sh


def attach(model, dist, group=None):
    """Switch a model (models.FM / FNN* / theano_fnn.FNN_wnzh) to data-parallel steps.  The reference's
    batch-size-dependent constants see the GLOBAL batch: 'mean' reduce and the decay 1-2*lambda*lr/b_size."""
    from .engine import BagEngine
    if isinstance(model.engine, BagEngine):
        raise NotImplementedError("the SNN embedding-bag step has no data-parallel form yet (its bias is a vector and every field of a "
                                  "sample shares one gradient row); run it on one GPU")
    dp = DataParallelStep(model.engine, dist, group)
    model.dp = dp
    # kernels AND collectives of a step replay as one CUDA graph (NCCL is capturable); DCTR_DP_GRAPH=0 issues them eagerly
    model.use_graph = os.environ.get("DCTR_DP_GRAPH", "1") != "0"
    model._graph, model._warm = None, 0
    model.global_N = model.N * dp.world
    if dp.is_fm:
        model.engine.train_step = dp.train_step_fm
    else:
        model.engine.train_step = dp.train_step
    return dp


def shutdown(dist, models=()):
    """Tear the process group down.  Captured data-parallel steps hold NCCL work: drop the graphs first, drain the
    device, meet at a barrier -- and do not let a stuck communicator teardown keep a finished job alive."""
    import gc
    import sys
    import threading
    for m in models:
        m._graph = None
    gc.collect()
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    sys.stderr.flush()
    timer = threading.Timer(20.0, lambda: os._exit(0))
    timer.daemon = True
    timer.start()
    dist.destroy_process_group()
    timer.cancel()
