"""Device-side step engine: owns the HBM-resident state (embedding table, flat MLP parameters, activation
and workspace buffers) for one static batch size and issues the libdctr kernels of one train / predict step
on the current CUDA stream.  torch is used for device memory, streams and CUDA-graph capture only.

Layout in HBM
  table   float32 [X, ld]      row = [w | v_1..v_r | 0 pad], ld = round_up(d, 4): a row is ld/4 128-bit loads
  params  float32 flat         [W_0 (K_0 x ldw_0) | b_0 | ... | w_out (K) | b_out | bias(w_0 / fm b)], 16 B aligned
  grads   float32 flat         same layout (one all-reduce buffer, one dense-SGD launch)
  ins[l]  float32 [N, ld_l]    input of layer l (ins[0] = act0(z)), ld_l = round_up(width, 4), pad columns 0
"""
import os

import torch

from . import _lib as L
from . import ops as _ops          # registers torch.ops.deepctr.* (the C-ABI of libdctr.so as PyTorch custom ops)

O = torch.ops.deepctr


def _r4(n):
    return (n + 3) // 4 * 4


class StepEngine:
    def __init__(self, X, F, rank, hidden, batch_size, cross="none", bias_first=False, act0="none", act="tanh",
                 scatter_mode="sum", device="cuda", ld=None, gemm=None):
        L.lib()  # fail loudly when libdctr.so is missing
        # MLP GEMM path: "tc" = tcgen05 3xTF32 (gemm_tc.cu), "simt" = fp32 FFMA (gemm.cu); "auto" keeps the SIMT kernel
        # below 1024 samples (fewer than 8 tensor-core row tiles)
        gemm = gemm or os.environ.get("DCTR_GEMM", "auto")
        assert gemm in ("auto", "tc", "simt"), gemm
        self.gemm_fn = "dctr_gemm_tc" if (gemm == "tc" or (gemm == "auto" and int(batch_size) >= 1024)) else "dctr_gemm"
        self.dev = torch.device(device)
        self.X, self.F, self.d, self.N = int(X), int(F), int(rank) + 1, int(batch_size)
        self.ld = ld or _r4(self.d)
        self.cross, self.bias_first, self.act0, self.act = cross, bool(bias_first), act0, act
        self.acts = None          # optional per-layer override: acts[l] = activation that produced ins[l], l >= 1 (FNN_wnzh.py:163)
        self.scatter_mode = L.SCATTER_SEQ_DECAY if scatter_mode == "seq_decay" else L.SCATTER_SUM
        F_, d_ = self.F, self.d
        self.crossw = {"none": 0, "ip": F_ * (F_ - 1) // 2, "op": d_ * d_}[cross]
        self.D0 = self._input_width()
        self.widths = [self.D0] + list(hidden)          # inputs of layers 0..L-1; the last layer maps to 1
        self.nl = len(self.widths)
        assert self.nl >= 2, "need at least one hidden layer"
        f32 = dict(dtype=torch.float32, device=self.dev)
        i32 = dict(dtype=torch.int32, device=self.dev)
        N = self.N
        self.table = torch.zeros(self.X, self.ld, **f32)
        # ---- flat parameter / gradient buffers ----
        self.w_off, self.b_off, self.ldw = [], [], []
        off = 0
        for l in range(self.nl):
            K = self.widths[l]
            Hout = self.widths[l + 1] if l + 1 < self.nl else 1
            ldw = _r4(Hout) if l + 1 < self.nl else 1
            self.ldw.append(ldw)
            self.w_off.append(off)
            off += _r4(K * ldw)
            self.b_off.append(off)
            off += _r4(Hout)
        self.bias_off = off
        off += self._bias_len()
        self.n_params = off
        self.params = torch.zeros(off, **f32)
        self.grads = torch.zeros(off + 4, **f32)          # + a 4-float tail: the loss rides along in the data-parallel all-reduce
        # ---- activations / scratch ----
        self.ldi = [_r4(w) for w in self.widths]
        self.ins = [torch.zeros(N, self.ldi[l], **f32) for l in range(self.nl)]
        maxw = max(self.ldi)
        self.dbuf = [torch.zeros(N, maxw, **f32) for _ in range(3)]
        self.dz0 = torch.zeros(N, self.ldi[0], **f32)         # d loss / d z of the embedding layer: its own, narrower buffer
        self.ids = torch.zeros(N, F_, **i32)
        self.wts = torch.ones(N, F_, **f32)
        self.y = torch.zeros(N, **f32)
        self.logit = torch.zeros(N, **f32)
        self.pred = torch.zeros(N, **f32)
        self.loss_n = torch.zeros(N, **f32)
        self.dlogit = torch.zeros(N, **f32)
        self.loss = torch.zeros(1, **f32)
        self.masks = [None] * self.nl                   # mask on ins[l]; tensors [N, ldi] or [1, ldi]
        M = N * F_
        self.M = M
        self.sorted_rows = torch.zeros(M, **i32)
        self.sorted_pos = torch.zeros(M, **i32)
        self.sort_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device=self.dev)
        self.scat_ws = torch.zeros(L.lib().dctr_scatter_workspace(M, self.ld), dtype=torch.uint8, device=self.dev)
        self.ldg = self.ld
        self.ge = torch.zeros(N, F_, self.ldg, **f32) if cross != "none" else None
        # split-K plan for the weight gradients: aim at ~2 waves of 64x64 tiles
        self.splits = []
        ws_elems = 4
        for l in range(self.nl - 1):
            K, H = self.widths[l], self.widths[l + 1]
            tiles = -(-K // 64) * -(-H // 64)
            s = max(1, min(max(1, N // 256), -(-2 * 148 // tiles)))
            self.splits.append(s)
            ws_elems = max(ws_elems, s * K * self.ldw[l])
        self.splitk_ws = torch.zeros(ws_elems, **f32)
        # workspace of the single-launch reductions (ticket + CTA partials): zero-filled once, left zeroed by the kernels
        self.col_ws = torch.zeros(L.lib().dctr_reduce_workspace(N, maxw + 4), dtype=torch.uint8, device=self.dev)
        self.col_ws2 = torch.zeros_like(self.col_ws)           # reductions issued from the weight-gradient stream
        self.side = torch.cuda.Stream(device=self.dev)
        self.side2 = torch.cuda.Stream(device=self.dev)
        self.overlap_wgrad = True
        self._wgrad_pending = False
        self.adam = None
        self.ftrl = None
        self.fuse_out_unit = True
        self._out_unit_done = False
        # programmatic dependent launch pays when the step is a chain of short kernels (measured: neutral at N = 4096 with the
        # weight-gradient stream, -3 % at N = 65536 where parked CTAs take SMs from the concurrent streams)
        self.launch_mode = 2 if self.N <= 8192 else 0
        self.overlap_sort = True
        self.wts_are_ones = True
        self.upd_rows = self.upd_pos = None     # use_update_list()
        self.dropout = None                     # (flat mask buffer, keep_prob, seed, device state): masks drawn inside the step
        # Small two-hidden-layer nets: the whole MLP forward + backward as ONE fp32 FFMA kernel per 32-sample tile and the weight
        # gradients (+ SGD) as one cluster kernel (csrc/mlp2.cu) instead of ~14 per-layer launches.  Opt-in (DCTR_MLP2=on):
        # measured at N = 4096, 177-300-100-1 it LOSES to the per-layer tcgen05 chain (0.264 vs 0.158 ms per step,
        # profiles/r02_bench_fnn_ipinyou_mlp2_v3.json) -- fp32 SIMT cannot match the tensor cores even at these sizes.
        # DCTR_MLP2=tc: the same fusion on the tensor cores (csrc/mlp2tc.cu: forward + data gradients of a 128-sample tile chained
        # through tcgen05 with the activations handed from epilogue to A tile in shared memory); weight gradients by
        # DCTR_MLP2_WGRAD = gemm (split-K TN GEMMs, one side stream per layer) | cluster (the FFMA cluster kernel of mlp2.cu).
        # Default (DCTR_MLP2=auto): the tcgen05 fusion wherever it applies and the engine is on the tensor-core path -- measured at
        # N = 4096, 177-300-100-1: 0.128 ms per step against 0.153 for the per-layer chain (profiles/r02_bench_fnn_ipinyou_mlp2tc.json).
        mode = os.environ.get("DCTR_MLP2", "auto")
        ok = self.nl == 3 and bool(L.lib().dctr_mlp2_supported(self.widths[0], self.widths[1], self.widths[2]))
        ok_tc = self.nl == 3 and bool(L.lib().dctr_mlp2tc_supported(self.widths[0], self.widths[1], self.widths[2]))
        self.mlp2_tc = ok_tc and (mode == "tc" or (mode == "auto" and self.gemm_fn == "dctr_gemm_tc"))
        self.mlp2_wgrad_gemm = self.mlp2_tc and os.environ.get("DCTR_MLP2_WGRAD", "gemm") == "gemm"
        self.fused_mlp2 = (ok and mode == "on") or (self.mlp2_tc and (ok or self.mlp2_wgrad_gemm))
        if self.mlp2_tc and not self.fused_mlp2:
            self.mlp2_tc = False
        self._mlp2_sgd = False                  # this step's fused kernels apply plain SGD to the MLP themselves
        self._bias_sgd_done = False
        self._mlp2_cfg = None                   # (lr, decay_hidden, decay_out, train_bias) when they do
        self._mlp2_tr_event = None
        if self.fused_mlp2 and not self.mlp2_wgrad_gemm:
            self.mlp2_ws = torch.zeros(L.lib().dctr_mlp2_workspace(N, self.widths[0], self.widths[1], self.widths[2]), dtype=torch.uint8, device=self.dev)
            # transposed copies of the two hidden-layer weights, rebuilt on the second stream at the start of every step: the
            # backward GEMMs then stream row-major [k][column] slabs like the forward ones
            self.W1t = torch.zeros(self.widths[1], self.ldi[0], **f32)
            self.W2t = torch.zeros(self.widths[2], self.ldi[1], **f32)
        if self.mlp2_tc:
            self.mlp2tc_ws = torch.zeros(L.lib().dctr_mlp2tc_workspace(N, self.widths[2]), dtype=torch.uint8, device=self.dev)
            # the hidden-layer weights as tf32 hi / lo tiles in the kernel's shared-memory layout and consumption order, rebuilt on
            # the second stream at the start of every step
            self.W12_packed = torch.zeros(L.lib().dctr_mlp2tc_packed_bytes(*self.widths), dtype=torch.uint8, device=self.dev)
            self.side3 = torch.cuda.Stream(device=self.dev)
            self.side4 = torch.cuda.Stream(device=self.dev)
            self.col_ws3 = torch.zeros_like(self.col_ws)
            self.splitk_ws3 = torch.zeros_like(self.splitk_ws)
        self.skip_row = None        # set when the weights are 0 / 1 and 0 only marks this (all-zero pad) row: no weight gather in the fold

    # ---- the embedding front end (overridden by BagEngine) -------------------------------------------
    def _input_width(self):
        return self.F * self.d + self.crossw + 1          # [e_0 .. e_{F-1} | cross | bias]

    def _bias_len(self):
        return 4                                          # w_0 / fm b: one float, padded

    def _embed_forward(self, ids, wts, table, X):
        O.gather_concat_fwd(ids.view(self.N, self.F), None if wts is None else wts.view(self.N, self.F), table[:X], self.d,
                            L.CROSS[self.cross], self.bias_first, self.bias(), L.ACT[self.act0], self.ins[0])

    # ---- parameter views -------------------------------------------------------------------------
    def W(self, l, buf=None):
        buf = self.params if buf is None else buf
        K = self.widths[l]
        return buf[self.w_off[l]:self.w_off[l] + K * self.ldw[l]].view(K, self.ldw[l])

    def b(self, l, buf=None):
        buf = self.params if buf is None else buf
        Hout = self.widths[l + 1] if l + 1 < self.nl else 1
        return buf[self.b_off[l]:self.b_off[l] + Hout]

    def bias(self, buf=None):
        buf = self.params if buf is None else buf
        return buf[self.bias_off:self.bias_off + 1]

    def set_layer(self, l, W, b):
        """W: [K, Hout] array-like, b: [Hout]."""
        W = torch.as_tensor(W, dtype=torch.float32).reshape(self.widths[l], -1)
        self.W(l)[:, :W.shape[1]] = W.to(self.dev)
        self.b(l).copy_(torch.as_tensor(b, dtype=torch.float32).reshape(-1))

    def get_layer(self, l):
        Hout = self.widths[l + 1] if l + 1 < self.nl else 1
        return self.W(l)[:, :Hout].cpu().numpy(), self.b(l).cpu().numpy()

    def set_table(self, table):
        t = torch.as_tensor(table, dtype=torch.float32)
        self.table.zero_()
        self.table[:t.shape[0], :t.shape[1]] = t.to(self.dev)      # a row shard may be one row short of the allocation

    def get_table(self):
        return self.table[:, :self.d].cpu().numpy()

    def act_of(self, l):
        """activation that produced ins[l]: act0 for the embedding layer, otherwise the hidden activation (or its override)"""
        if l == 0:
            return self.act0
        return self.acts[l] if self.acts is not None else self.act

    def use_update_list(self, J):
        """Sparse update over an explicit (row, slot) list of J entries per sample instead of the [N, F] id block: the Theano
        script updates EVERY feature of a sample, also one whose slot a later feature of the same field took
        (FNN_wnzh.py:300-306).  dctr_arrange_update_list fills upd_rows / upd_pos, dctr_sort_pairs orders them."""
        i32 = dict(dtype=torch.int32, device=self.dev)
        self.J = int(J)
        self.M = self.N * self.J
        self.upd_rows = torch.zeros(self.M, **i32)
        self.upd_pos = torch.zeros(self.M, **i32)
        self.sorted_rows = torch.zeros(self.M, **i32)
        self.sorted_pos = torch.zeros(self.M, **i32)
        self.sort_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(self.M), dtype=torch.uint8, device=self.dev)
        self.scat_ws = torch.zeros(L.lib().dctr_scatter_workspace(self.M, self.ld), dtype=torch.uint8, device=self.dev)

    # ---- kernels ---------------------------------------------------------------------------------
    def _st(self):
        return torch.cuda.current_stream(self.dev).cuda_stream

    def _gemm(self, *a):
        (O.gemm_tc if self.gemm_fn == "dctr_gemm_tc" else O.gemm)(*a)

    def forward(self, with_loss, gscale=1.0, lookup=None):
        """lookup = (ids, rows, n_rows): gather from an already exchanged row buffer instead of the local table
        (row-sharded mode, dist.ShardedStep)."""
        N = self.N
        wts = None if self.wts_are_ones else self.wts
        ids, table, X = (self.ids, self.table, self.X) if lookup is None else lookup
        self._embed_forward(ids, wts, table, X)
        if with_loss and self.fused_mlp2:
            if self.dropout is not None:
                flat, keep, seed, state = self.dropout
                O.dropout_masks(flat, keep, seed, state, None)      # the fused kernel applies the first mask itself
            self._mlp2_forward_backward(gscale)
            return
        if with_loss and self.dropout is not None:
            flat, keep, seed, state = self.dropout      # this step's tf.nn.dropout masks, the first one applied to act(z1) (FNN.py:87)
            O.dropout_masks(flat, keep, seed, state, self.ins[0])
        elif self.masks[0] is not None:
            self.ins[0].mul_(self.masks[0])     # injected masks (parity runs): plain elementwise product
        for l in range(self.nl - 1):
            self._gemm(L.GEMM_NN, N, self.widths[l + 1], self.widths[l], self.ins[l], self.W(l), self.ins[l + 1], L.EPI_FWD,
                       L.ACT[self.act_of(l + 1)], self.b(l), None, self.masks[l + 1], 1)
        last = self.nl - 1
        self._out_unit_done = False
        if with_loss and self.widths[last] <= 1024 and self.fuse_out_unit:
            # train step: the output unit forward AND backward in one pass over the last activation
            O.out_unit_train(self.ins[last], self.widths[last], self.W(last), self.b(last), self.y, float(gscale),
                             L.ACT[self.act_of(last)], self.masks[last], self.logit, self.pred, self.loss_n, self.dlogit,
                             float(gscale), self.loss, self.dbuf[0], self.W(last, self.grads), self.b(last, self.grads), self.col_ws)
            self._out_unit_done = True
            return
        # logit / pred / per-sample loss / dlogit and the reduced loss in one launch
        O.out_xent_fwd(self.ins[last], self.widths[last], self.W(last), self.b(last), self.y if with_loss else None, float(gscale),
                       self.logit, self.pred, self.loss_n, self.dlogit, float(gscale), self.loss if with_loss else None, self.col_ws)

    def _mlp2_forward_backward(self, gscale):
        """csrc/mlp2.cu, kernel 1: x -> a1 -> a2 -> output unit -> dz2 -> dz1 -> dz0 for every 32-sample tile; with plain SGD
        the kernel's last CTA also updates the output unit."""
        D0, H1, H2 = self.widths
        sgd = self._mlp2_cfg is not None
        lr, _, dout = self._mlp2_cfg[:3] if sgd else (0.0, 1.0, 1.0)
        self._mlp2_sgd = sgd
        main = torch.cuda.current_stream(self.dev)
        if self._mlp2_tr_event is not None:             # the transposes were issued on the second stream (train_step)
            main.wait_event(self._mlp2_tr_event)
            self._mlp2_tr_event = None
        else:
            self._mlp2_transpose()
        if self.mlp2_tc:
            O.mlp2tc_train(D0, H1, H2, self.ins[0], self.masks[0], L.ACT[self.act_of(0)], self.W12_packed, self.b(0), L.ACT[self.act_of(1)],
                           self.masks[1], self.b(1), L.ACT[self.act_of(2)], self.masks[2], self.W(2).view(-1),
                           self.b(2), self.y, float(gscale), float(gscale), self.ins[1], self.dbuf[0], self.dbuf[1], self.dz0, self.logit,
                           self.pred, self.loss_n, self.dlogit, self.loss, self.W(2, self.grads).view(-1), self.b(2, self.grads), sgd,
                           float(lr), float(dout), self.mlp2tc_ws)
            return
        O.mlp2_train(D0, H1, H2, self.ins[0], self.masks[0], L.ACT[self.act_of(0)], self.W(0), self.b(0), L.ACT[self.act_of(1)],
                     self.masks[1], self.W(1), self.b(1), L.ACT[self.act_of(2)], self.masks[2], self.W1t, self.W2t, self.W(2).view(-1),
                     self.b(2), self.y, float(gscale), float(gscale), self.ins[1], self.dbuf[0], self.dbuf[1], self.dz0, self.logit,
                     self.pred, self.loss_n, self.dlogit, self.loss, self.W(2, self.grads).view(-1), self.b(2, self.grads), sgd,
                     float(lr), float(dout), self.mlp2_ws)

    def _mlp2_transpose(self):
        D0, H1, H2 = self.widths
        if self.mlp2_tc:
            O.mlp2tc_pack(D0, H1, H2, self.W(0), self.W(1), self.W12_packed)
            return
        O.mlp2_transpose(D0, H1, H2, self.W(0), self.W(1), self.W1t, self.W2t)

    def _mlp2_wgrad(self):
        """kernel 2: both weight gradients and bias gradients as one launch of 8-CTA clusters (gradients into self.grads, or
        SGD in place); runs on the second stream next to the sparse update, which only needs dz0."""
        D0, H1, H2 = self.widths
        sgd = self._mlp2_sgd
        lr, dhid, _ = self._mlp2_cfg[:3] if sgd else (0.0, 1.0, 1.0)
        ov = self.overlap_wgrad
        main = torch.cuda.current_stream(self.dev)
        if self.mlp2_wgrad_gemm:
            # dz1 = dbuf[0] pairs with ins[0], dz2 = dbuf[1] with ins[1].  Per layer: split-K TN GEMM -> fold of the slabs, and the
            # bias column sums; the layers are independent of each other and the bias sums of the GEMMs, so there are three side
            # streams (all start when the fused kernel ends).  Plain SGD on one GPU: the fold and the column sums apply the update
            # themselves (W = W * decay - lr * gW), the gradients never go to memory and there is no dense sweep afterwards.
            def on(side):
                if ov:
                    side.wait_stream(main)
                return torch.cuda.stream(side if ov else main)
            for l, side, sws in ((0, self.side2, self.splitk_ws), (1, self.side3, self.splitk_ws3)):
                with on(side):
                    K, Hl, s_ = self.widths[l], self.widths[l + 1], self.splits[l]
                    dpre, gW = self.dbuf[l], self.W(l, self.grads)
                    if s_ > 1:
                        slabs = sws[:s_ * K * self.ldw[l]].view(s_, K, self.ldw[l])
                        self._gemm(L.GEMM_TN, K, Hl, self.N, self.ins[l], dpre, slabs, L.EPI_NONE, 0, None, None, None, s_)
                        if sgd:
                            O.splitk_reduce_sgd(slabs, s_, K, Hl, self.ldw[l], self.W(l), float(lr), float(dhid))
                        else:
                            O.splitk_reduce(slabs, s_, K, Hl, self.ldw[l], gW)
                    else:
                        self._gemm(L.GEMM_TN, K, Hl, self.N, self.ins[l], dpre, gW, L.EPI_NONE, 0, None, None, None, 1)
                        if sgd:
                            lo = self.w_off[l]
                            O.dense_sgd(self.params[lo:lo + K * self.ldw[l]], self.grads[lo:lo + K * self.ldw[l]], K * self.ldw[l], float(lr), float(dhid))
            with on(self.side4):
                for l, cws in ((0, self.col_ws2), (1, self.col_ws3)):
                    cw = cws if ov else self.col_ws
                    if sgd:
                        O.col_wsum_sgd(self.dbuf[l], self.widths[l + 1], None, self.b(l), float(lr), float(dhid), cw)
                    else:
                        O.col_wsum(self.dbuf[l], self.widths[l + 1], None, self.b(l, self.grads), cw)
            if ov:
                self._wgrad_pending = True
            return
        if ov:
            self.side2.wait_stream(main)
        with torch.cuda.stream(self.side2 if ov else main):
            O.mlp2_wgrad(D0, H1, H2, self.ins[0], self.ins[1], self.dbuf[0], self.dbuf[1], self.W(0), self.b(0), self.W(1), self.b(1),
                         self.W(0, self.grads), self.b(0, self.grads), self.W(1, self.grads), self.b(1, self.grads), sgd, float(lr),
                         float(dhid), self.mlp2_ws)
        if ov:
            self._wgrad_pending = True

    def backward(self):
        """Fills self.grads (dense) and leaves d loss / d z in self.dbuf[k]; returns that buffer."""
        if self.fused_mlp2:
            self._mlp2_wgrad()
            return self.dz0
        N = self.N
        last = self.nl - 1
        H = self.widths[last]
        cur = 0
        # d(last hidden), d w_out and d b_out in one launch (unless the forward already produced them)
        if not self._out_unit_done:
            O.out_bwd_fused(self.dlogit, self.W(last), self.ins[last], H, L.ACT[self.act_of(last)], self.masks[last], self.dbuf[cur],
                            self.W(last, self.grads), self.b(last, self.grads), self.col_ws)
        self._out_unit_done = False
        # Per layer, the weight/bias gradients (TN GEMM [+ split-K fold], column sum) only need dz_l, as does the data
        # gradient that continues the chain: they run on a second stream next to it (three rotating dz buffers, so the
        # chain is two layers ahead before it overwrites what a weight-gradient GEMM still reads).
        ov = self.overlap_wgrad
        main = torch.cuda.current_stream(self.dev)
        readers = [None, None, None]
        for l in range(last - 1, -1, -1):
            K, Hl = self.widths[l], self.widths[l + 1]
            dpre = self.dbuf[cur]
            s = self.splits[l]
            gW = self.W(l, self.grads)
            if ov:
                self.side2.wait_stream(main)
            with torch.cuda.stream(self.side2 if ov else main):
                cws = self.col_ws2 if ov else self.col_ws
                if s > 1:
                    slabs = self.splitk_ws[:s * K * self.ldw[l]].view(s, K, self.ldw[l])
                    self._gemm(L.GEMM_TN, K, Hl, N, self.ins[l], dpre, slabs, L.EPI_NONE, 0, None, None, None, s)
                    O.splitk_reduce(slabs, s, K, Hl, self.ldw[l], gW)
                else:
                    self._gemm(L.GEMM_TN, K, Hl, N, self.ins[l], dpre, gW, L.EPI_NONE, 0, None, None, None, 1)
                O.col_wsum(dpre, Hl, None, self.b(l, self.grads), cws)
                if ov:
                    readers[cur] = self.side2.record_event()
                    self._wgrad_pending = True
            nxt = (cur + 1) % 3
            if readers[nxt] is not None:
                main.wait_event(readers[nxt])
            dout = self.dbuf[nxt] if l > 0 else self.dz0
            self._gemm(L.GEMM_NT, N, K, Hl, dpre, self.W(l), dout, L.EPI_DGRAD, L.ACT[self.act_of(l)], None, self.ins[l],
                       self.masks[l], 1)
            cur = nxt
        if last > 0:
            return self.dz0
        return self.dbuf[cur]

    def sort_ids(self, ids=None, M=None, rows=None, pos=None, ws=None):
        payload = None
        if ids is None:
            ids, payload = (self.ids, None) if self.upd_rows is None else (self.upd_rows, self.upd_pos)
        M = self.M if M is None else M
        O.sort_pairs(ids, payload, M, self.X, self.sorted_rows if rows is None else rows, self.sorted_pos if pos is None else pos,
                     self.sort_ws if ws is None else ws)

    def embedding_update(self, dz, lr, decay=1.0, train_bias=True):
        """dz: [N, ldd] gradient w.r.t. z.  Requires sort_ids() to have run on this stream (or joined)."""
        F, d = self.F, self.d
        wts = None if self.wts_are_ones else self.wts
        if train_bias:
            col = 0 if self.bias_first else F * d + self.crossw
            if self._mlp2_sgd:          # plain SGD with the fused MLP kernels: the bias takes its step right here
                O.col_wsum_sgd(dz[:, col:col + 1], 1, None, self.bias(), float(lr), 1.0, self.col_ws)
                self._bias_sgd_done = True
            else:
                O.col_wsum(dz[:, col:col + 1], 1, None, self.bias(self.grads), self.col_ws)
        if self.cross != "none":
            O.cross_bwd(self.ids.view(self.N, F), None if wts is None else wts.view(self.N, F), self.table, d, L.CROSS[self.cross], dz, self.ge)
            g, g_ld_f, w = self.ge, self.ldg, None
        else:
            g, g_ld_f, w = dz[:, (1 if self.bias_first else 0):], d, wts
        self.table_update(self.sorted_rows, self.sorted_pos, self.M, F, g, g_ld_f, w, lr, decay, self.scat_ws)

    def table_update(self, rows, pos, M, F, g, g_ld_f, w, lr, decay, ws):
        """Fold the sorted (row, position) list into the table: SGD in place, or -- Adam / FTRL -- into the dense gradient
        buffer followed by the dense sweep.  g: [N, >= F * g_ld_f] gradient rows (field f at column f * g_ld_f) or a per-slot
        buffer [N, F, ldg].  Also the entry point of the multi-GPU steps (global / received lists)."""
        if M == 0:
            rows = pos = self.sorted_rows            # never read
        if self.dense_optimizer:
            if self.adam is not None:
                O.adam_tick(self.adam_state, float(lr), self.adam[0], self.adam[1])
            O.scatter_sgd(self.scatter_mode, rows, pos, M, F, self.d, g, g_ld_f, w, -1, -1.0, 1.0, self.table_g, ws)
            if self.adam is not None:
                b1, b2, eps = self.adam
                O.dense_adam(self.table, self.table_g, self.table_m, self.table_v, self.X * self.ld, self.adam_state, b1, b2, eps, 0.0, True)
            else:
                lr_power, l1, l2 = self.ftrl
                O.dense_ftrl(self.table, self.table_g, self.table_n, self.table_z, self.X * self.ld, float(lr), lr_power, l1, l2, 0.0, True)
            return
        if self.skip_row is not None:
            O.scatter_sgd(self.scatter_mode, rows, pos, M, F, self.d, g, g_ld_f, None, int(self.skip_row), float(lr), float(decay),
                          self.table, ws)
            return
        O.scatter_sgd(self.scatter_mode, rows, pos, M, F, self.d, g, g_ld_f, w, -1, float(lr), float(decay), self.table, ws)

    def enable_adam(self, eps=1e-8, beta1=0.9, beta2=0.999):
        """Switch the step to TensorFlow's Adam over the table AND the MLP (tf_util.py:18-21).  TF densifies the table gradient
        (FNN.py:46-49), so the slots of every row decay every step: the fold writes the summed per-row gradients into a
        dense buffer G (dctr_scatter_sgd with table = G, lr = -1, decay = 1: G[row] = sum g) and one streaming
        dctr_dense_adam pass over table | G | m | v applies the update and clears G (7 x 4 x X x ld bytes per step)."""
        assert self.scatter_mode == L.SCATTER_SUM, "Adam belongs to the TF-form models (SUM scatter)"
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.adam = (float(beta1), float(beta2), float(eps))
        self.adam_state = torch.zeros(4, **f32)                 # [t, lr_t, beta1^t, beta2^t], advanced on the device
        self.table_g, self.table_m, self.table_v = (torch.zeros_like(self.table) for _ in range(3))
        self.params_m, self.params_v = torch.zeros_like(self.grads), torch.zeros_like(self.grads)

    def enable_ftrl(self, init_accum=0.1, lr_power=-0.5, l1=0.0, l2=0.0):
        """Switch the step to TensorFlow's FtrlOptimizer(lr) over the table AND the MLP (tf_util.py:22-24; TF defaults).  Dense
        for the same reason as Adam: the fold writes per-row gradient sums into G, one streaming dctr_dense_ftrl pass over
        table | G | accum | linear applies the update and clears G."""
        assert self.scatter_mode == L.SCATTER_SUM, "FTRL belongs to the TF-form models (SUM scatter)"
        self.ftrl = (float(lr_power), float(l1), float(l2))
        self.table_g, self.table_z = torch.zeros_like(self.table), torch.zeros_like(self.table)
        self.table_n = torch.full_like(self.table, float(init_accum))
        self.params_n, self.params_z = torch.full_like(self.grads, float(init_accum)), torch.zeros_like(self.grads)

    @property
    def dense_optimizer(self):
        """True when the table update is a dense sweep (Adam / FTRL): every row moves even without a gradient."""
        return self.adam is not None or self.ftrl is not None

    def join_wgrad(self):
        """self.grads is complete on the current stream after this (weight gradients run on a side stream)."""
        if self._wgrad_pending:
            torch.cuda.current_stream(self.dev).wait_stream(self.side2)
            if self.mlp2_wgrad_gemm:
                torch.cuda.current_stream(self.dev).wait_stream(self.side3)
                torch.cuda.current_stream(self.dev).wait_stream(self.side4)
            self._wgrad_pending = False

    def dense_update(self, lr, last_decay=1.0, mlp_decay=1.0):
        """last_decay: L2 on the output unit only (FNN_wnzh.py:173); mlp_decay: L2 on every MLP weight and bias but not on the
        embedding-side bias (SNN_DAE.py:151-153)."""
        self.join_wgrad()
        P, G, n = self.params, self.grads, self.n_params

        def sgd(lo, hi, decay):
            O.dense_sgd(P[lo:hi], G[lo:hi], hi - lo, float(lr), float(decay))
        if self._mlp2_sgd:                 # the fused kernels updated the MLP; what is left is the embedding-side bias
            self._mlp2_sgd = False
            if self._mlp2_cfg[3] and not self._bias_sgd_done:
                sgd(self.bias_off, n, 1.0)
            self._bias_sgd_done = False
            return
        if self.adam is not None:          # same step counter as the table (embedding_update ticked it)
            b1, b2, eps = self.adam
            O.dense_adam(P, G, self.params_m, self.params_v, n, self.adam_state, b1, b2, eps, 0.0, False)
            return
        if self.ftrl is not None:
            lr_power, l1, l2 = self.ftrl
            O.dense_ftrl(P, G, self.params_n, self.params_z, n, float(lr), lr_power, l1, l2, 0.0, False)
            return
        if mlp_decay != 1.0:
            sgd(0, self.bias_off, mlp_decay)
            sgd(self.bias_off, n, 1.0)
        elif last_decay == 1.0:
            sgd(0, n, 1.0)
        else:   # Theano: cost += lambda1*(|w3|^2 + b3^2) touches the output unit only (FNN_wnzh.py:173)
            lo, hi = self.w_off[self.nl - 1], self.bias_off
            sgd(0, lo, 1.0)
            sgd(lo, hi, last_decay)
            sgd(hi, n, 1.0)

    def train_step(self, lr, decay=1.0, gscale=1.0, train_bias=True, last_decay=1.0, mlp_decay=1.0):
        """One full step on device buffers ids/wts/y (+ masks).  The row sort only depends on the ids, so it
        runs on a side stream concurrently with the MLP forward/backward."""
        L.call("dctr_set_launch_mode", self.launch_mode)
        cur = torch.cuda.current_stream(self.dev)
        if self.overlap_sort:
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                self.sort_ids()
        else:                                    # per-kernel timing (bench.py): everything on one stream
            self.sort_ids()
        # plain SGD on one GPU: the fused MLP kernels apply the update themselves (the multi-GPU steps call forward / backward
        # directly and all-reduce self.grads first, Adam / FTRL need the gradients)
        plain = self.fused_mlp2 and not self.dense_optimizer
        if self.fused_mlp2 and self.overlap_wgrad:     # W^T for the backward GEMMs: off the critical path, next to the gather
            self.side2.wait_stream(cur)
            with torch.cuda.stream(self.side2):
                self._mlp2_transpose()
                self._mlp2_tr_event = self.side2.record_event()
        self._mlp2_cfg = (float(lr), float(mlp_decay), float(last_decay if mlp_decay == 1.0 else mlp_decay), bool(train_bias)) if plain else None
        self.forward(True, gscale)
        dz = self.backward()
        if self.overlap_sort:
            cur.wait_stream(self.side)
        self.embedding_update(dz, lr, decay, train_bias)
        self.dense_update(lr, last_decay, mlp_decay)
        self._mlp2_cfg = None


class BagEngine(StepEngine):
    """Embedding-BAG front end: ins[0] = sigmoid(sum_f table[ids[n,f]] + b0) with H0-wide rows instead of the
    per-field concat -- the first layer of the SNN models (get_fi_h1_y, SNN_DAE.py:234-260).  Everything behind it is the
    StepEngine: the first dgrad GEMM's epilogue applies sigmoid' and leaves gx*x*(1-x) in dz0, the sorted fold adds that
    sample row into every row the sample touched (g_ld_f = 0) and its column sums are b0's gradient (SNN_DAE.py:282-288)."""

    def __init__(self, X, F, H0, hidden, batch_size, act="tanh", device="cuda", gemm=None):
        StepEngine.__init__(self, X, F, H0 - 1, hidden, batch_size, cross="none", bias_first=False, act0="sigmoid", act=act,
                            scatter_mode="sum", device=device, gemm=gemm)

    def _input_width(self):
        return self.d                                     # d = H0: the bag output IS the MLP input

    def _bias_len(self):
        return _r4(self.d)                                # bb0 [H0]

    def bias(self, buf=None):
        buf = self.params if buf is None else buf
        return buf[self.bias_off:self.bias_off + self.d]

    def _embed_forward(self, ids, wts, table, X):
        O.bag_fwd(ids.view(self.N, self.F), None if wts is None else wts.view(self.N, self.F), table[:X], self.d, self.bias(),
                  L.ACT[self.act0], self.ins[0])

    def embedding_update(self, dz, lr, decay=1.0, train_bias=True):
        if train_bias:          # bb0 -= lr * sum_t gx_t x_t (1 - x_t)   (SNN_DAE.py:284)
            O.col_wsum(dz, self.d, None, self.bias(self.grads), self.col_ws)
        wts = None if self.wts_are_ones else self.wts
        self.table_update(self.sorted_rows, self.sorted_pos, self.M, self.F, dz, 0, wts, lr, decay, self.scat_ws)


class FMEngine:
    """FM (and LR with rank 0): fm_fwd -> [sort on a side stream] -> scatter_sgd_fm -> bias SGD (+ L2 sweep)."""

    def __init__(self, X, F, rank, batch_size, device="cuda", ld=None):
        L.lib()
        self.dev = torch.device(device)
        self.X, self.F, self.d, self.N = int(X), int(F), int(rank) + 1, int(batch_size)
        self.ld = ld or _r4(self.d)
        self.lds = _r4(self.d)
        f32 = dict(dtype=torch.float32, device=self.dev)
        i32 = dict(dtype=torch.int32, device=self.dev)
        N, M = self.N, self.N * self.F
        self.M = M
        self.table = torch.zeros(self.X, self.ld, **f32)
        self.params = torch.zeros(4, **f32)     # [b, -, -, -]
        self.grads = torch.zeros(8, **f32)          # [gb, -, -, -, loss slot of the data-parallel all-reduce, ...]
        self.n_params = 4
        self.ids = torch.zeros(N, self.F, **i32)
        self.wts = torch.ones(N, self.F, **f32)
        self.y = torch.zeros(N, **f32)
        self.logit = torch.zeros(N, **f32)
        self.pred = torch.zeros(N, **f32)
        self.loss_n = torch.zeros(N, **f32)
        self.g = torch.zeros(N, **f32)
        self.s = torch.zeros(N, self.lds, **f32)
        self.loss = torch.zeros(1, **f32)
        self.sorted_rows = torch.zeros(M, **i32)
        self.sorted_pos = torch.zeros(M, **i32)
        self.sort_ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device=self.dev)
        self.scat_ws = torch.zeros(L.lib().dctr_scatter_workspace(M, self.ld), dtype=torch.uint8, device=self.dev)
        self.side = torch.cuda.Stream(device=self.dev)
        self.overlap_sort = True
        self.adam = None
        self.ftrl = None
        self.l2_ws = torch.zeros(L.lib().dctr_reduce_workspace(1, 1), dtype=torch.uint8, device=self.dev)
        self.launch_mode = 1 if self.N <= 16384 else 0       # FM step = 10 short kernels: -9 % with dependent launches
        self.wts_are_ones = False

    def bias(self, buf=None):
        return (self.params if buf is None else buf)[0:1]

    def set_table(self, table):
        t = torch.as_tensor(table, dtype=torch.float32)
        self.table.zero_()
        self.table[:, :t.shape[1]] = t.to(self.dev)

    def get_table(self):
        return self.table[:, :self.d].cpu().numpy()

    def _st(self):
        return torch.cuda.current_stream(self.dev).cuda_stream

    def forward(self, with_loss, gscale=1.0, lam=0.0):
        wts = None if self.wts_are_ones else self.wts.view(self.N, self.F)
        O.fm_fwd(self.ids.view(self.N, self.F), wts, self.table, self.d, self.bias(), self.y if with_loss else None, float(gscale),
                 self.logit, self.pred, self.loss_n, self.g, self.s if with_loss else None)
        if with_loss:
            O.vec_sum(self.loss_n, self.N, float(gscale), self.loss)
            if lam != 0.0:      # the loss the reference fetches carries lam * (l2_loss(W) + l2_loss(V) + l2_loss(b)) (FM.py:39, LR.py:37)
                O.l2_loss_add(self.table.view(-1), self.bias(), float(lam), self.loss, self.l2_ws)

    def enable_adam(self, eps=1e-8, beta1=0.9, beta2=0.999):
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.adam = (float(beta1), float(beta2), float(eps))
        self.adam_state = torch.zeros(4, **f32)
        self.table_g, self.table_m, self.table_v = (torch.zeros_like(self.table) for _ in range(3))
        self.params_m, self.params_v = torch.zeros_like(self.grads), torch.zeros_like(self.grads)

    def enable_ftrl(self, init_accum=0.1, lr_power=-0.5, l1=0.0, l2=0.0):
        self.ftrl = (float(lr_power), float(l1), float(l2))
        self.table_g, self.table_z = torch.zeros_like(self.table), torch.zeros_like(self.table)
        self.table_n = torch.full_like(self.table, float(init_accum))
        self.params_n, self.params_z = torch.full_like(self.grads, float(init_accum)), torch.zeros_like(self.grads)

    @property
    def dense_optimizer(self):
        return self.adam is not None or self.ftrl is not None

    def sort_ids(self):
        O.sort_pairs(self.ids, None, self.M, self.X, self.sorted_rows, self.sorted_pos, self.sort_ws)

    def train_step(self, lr, lam=0.0, gscale=1.0, dense_l2=True):
        L.call("dctr_set_launch_mode", self.launch_mode)
        cur = torch.cuda.current_stream(self.dev)
        if self.overlap_sort:
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                self.sort_ids()
        else:                                    # per-kernel timing (bench.py): everything on one stream
            self.sort_ids()
        self.forward(True, gscale, lam)
        O.vec_sum(self.g, self.N, 1.0, self.grads)           # d loss / d b = sum_n g_n
        self.bias_update(lr, lam)          # before the join: the one-float update runs while the sort is still going
        if self.overlap_sort:
            cur.wait_stream(self.side)
        wts = None if self.wts_are_ones else self.wts
        self.apply_update(self.sorted_rows, self.sorted_pos, self.M, self.g, self.s, wts, lr, lam, dense_l2, self.scat_ws)

    def bias_update(self, lr, lam):
        """b from grads[0] = sum_n g_n (already all-reduced in the data-parallel step); ticks Adam's step counter."""
        P, G = self.params[:4], self.grads[:4]
        if self.adam is not None:
            b1, b2, eps = self.adam
            O.adam_tick(self.adam_state, float(lr), b1, b2)
            O.dense_adam(P, G, self.params_m[:4], self.params_v[:4], 4, self.adam_state, b1, b2, eps, float(lam), False)
        elif self.ftrl is not None:
            lr_power, l1, l2 = self.ftrl
            O.dense_ftrl(P, G, self.params_n[:4], self.params_z[:4], 4, float(lr), lr_power, l1, l2, float(lam), False)
        else:
            O.dense_sgd(P, G, 1, float(lr), float(1.0 - lr * lam))

    def apply_update(self, rows, pos, M, g, s, wts, lr, lam, dense_l2, ws):
        """Table update from a sorted (row, position) list: this rank's own, or the global one of the data-parallel step
        (bias_update must have run first: it advances Adam's step).  rows / pos / g / s / wts / ws are device tensors."""
        decay = 1.0 - lr * lam
        F, d, n = self.F, self.d, self.X * self.ld
        if self.dense_optimizer:
            # TF Adam / FTRL are dense (tf_util.py:18-24): per-row gradients into G, then one sweep over table | G | slots with the
            # L2 term's gradient lam * theta folded in (FM.py:39, LR.py:37); the bias the same way
            O.scatter_grad_fm(rows, pos, M, F, d, g, s, wts, self.table, self.table_g, ws)
            if self.adam is not None:
                b1, b2, eps = self.adam
                O.dense_adam(self.table, self.table_g, self.table_m, self.table_v, n, self.adam_state, b1, b2, eps, float(lam), True)
            else:
                lr_power, l1, l2 = self.ftrl
                O.dense_ftrl(self.table, self.table_g, self.table_n, self.table_z, n, float(lr), lr_power, l1, l2, float(lam), True)
            return
        O.scatter_sgd_fm(rows, pos, M, F, d, g, s, wts, float(lr), float(decay), self.table, ws)
        if lam != 0.0 and dense_l2:   # FM.py:39: the L2 term reaches every row, touched or not
            O.decay_sweep(self.table, float(decay), rows, M)
