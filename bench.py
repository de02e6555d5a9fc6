#!/usr/bin/env python
"""bench.py -- one JSON line per run (see the driver contract in DESIGN.md section "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload fnn_ipinyou] [--impl reference]

A "step" is one full train step of the hot path (gather -> MLP fwd/bwd -> sorted segmented scatter-SGD ->
dense SGD) over one synthetic batch.  `value` is device-timed with the batch already resident in HBM;
`e2e` goes through the reference-facing model class with HOST numpy buffers (H2D + D2H inside).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (form, F, X, rank, hidden, cross, batch per GPU)
    "fnn_ipinyou": ("theano", 16, 1_000_000, 10, [300, 100], "none", 4096),
    "fm_ipinyou": ("fm", 16, 1_000_000, 10, [], "none", 4096),
    "fnn_ip_l3_ipinyou": ("tf", 16, 1_000_000, 10, [400, 400, 200], "ip", 4096),
    "fnn_op_l3_ipinyou": ("tf", 16, 1_000_000, 10, [800, 400, 200], "op", 4096),
    "fnn_ip_l7_ipinyou": ("tf", 16, 1_000_000, 10, [1000, 800, 600, 400, 200, 100, 50], "ip", 65536),
    "fnn_criteo": ("tf", 39, 1_636_608, 10, [400, 400], "none", 65536),
    # the optimizer baseline.py:93 actually runs on Criteo: adam, lr 1e-4, eps 1e-8, 'sum' (TF's dense Adam over table + MLP)
    "fnn_criteo_adam": ("tf", 39, 1_636_608, 10, [400, 400], "none", 65536),
    # BASELINE config 5: 26 categorical fields over a 10^8-row table, rows sharded over the ranks (row % world)
    "fnn_criteo_sharded": ("tf", 26, 100_000_000, 10, [400, 400], "none", 65536),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tensor=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons sampled while the timed region runs (NVML, else nvidia-smi every 200 ms)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.first = index, [], False, 0

    def begin(self):
        """the timed region starts here: earlier samples (NVML start-up, warm-up) are not reported"""
        self.first = max(0, len(self.rows) - 1)

    def _nvml_loop(self):
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        bits = [0x8, 0x40, 0x20, 0x4]        # hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap
        while not self.stop_flag:
            r = get_reasons(h)
            self.rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(mx)] +
                             ["Active" if r & b else "Not Active" for b in bits])
            time.sleep(0.01)

    def run(self):
        try:
            self._nvml_loop()              # NVML in-process: a sample every 10 ms
            return
        except Exception:
            pass
        while not self.stop_flag:          # fallback: the nvidia-smi query of the profiling recipe
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        self.rows = self.rows[self.first:]
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# synthetic workload (seeded; iPinYou-shaped: 16 fields, field cardinalities summing to X, Zipf(1.1) ids)
# ---------------------------------------------------------------------------------------------------
def make_pool(form, F, X, N, n_batches, seed):
    from deepctr_b200 import data
    rng = np.random.default_rng(seed)
    if F == 39:
        sizes = data.scaled_criteo_cat_sizes(X) if X != 13 + int(data.CRITEO_CAT_SIZES.sum()) else data.CRITEO_CAT_SIZES
        pool = []
        for _ in range(n_batches):
            ids, wts = data.make_criteo_batch(rng, N, sizes)
            pool.append((ids, wts, (rng.random(N) < 0.05).astype(np.float32)))
        return pool, sizes
    sizes = data.scaled_criteo_cat_sizes(X + 13) if F == 26 else data.ipinyou_field_sizes(X)
    pool = []
    for _ in range(n_batches):
        ids = data.make_field_batch(rng, N, sizes, 1.05 if F == 26 else 1.1)
        pool.append((ids, np.ones((N, F), np.float32), (rng.random(N) < 0.05).astype(np.float32)))
    return pool, sizes


def build_model(wl, N, seed=1234, shard=None):
    """Returns (model, train_fn(host ids, wts, y) -> (loss, preds)) through the reference-facing classes."""
    form, F, X, rank, hidden, cross, _ = WORKLOADS[wl]
    rng = np.random.default_rng(seed)
    d = rank + 1
    if form == "theano":
        from deepctr_b200.theano_fnn import FNN_wnzh
        from deepctr_b200 import data
        sizes = data.ipinyou_field_sizes(X)
        feat_field = np.repeat(np.arange(F), sizes)
        table = rng.uniform(-0.01, 0.01, (X, d)).astype(np.float32)
        m = FNN_wnzh(dict(w_0=0.0, table=table, feat_field=feat_field), batch_size=N, lr=0.001, dropout=0.5,
                     hidden1=hidden[0], hidden2=hidden[1], lambda1=0.0, lambda_fm=0.1)
        m.engine.W(2).uniform_(-0.05, 0.05)      # w3 = 0 (FNN_wnzh.py:130) would make step-1 gradients trivially 0
        return m, lambda ids, wts, y: m.train(ids, y)
    if form == "fm":
        from deepctr_b200.models import FM
        m = FM(N, [X, F, rank], ["uniform", -0.01, 0.01, [0x3210, 0x7654], None], ["sgd", 1e-3, "sum"], [0.0], "train", N)
        return m, lambda ids, wts, y: m.train_step({m.sp_id_hldr: ids.reshape(-1), m.sp_wt_hldr: wts.reshape(-1), m.lbl_hldr: y})
    import deepctr_b200.models as M
    from deepctr_b200 import data
    cls = {("none", 2): M.FNN, ("ip", 3): M.FNN_IP_L3, ("op", 3): M.FNN_OP_L3, ("ip", 5): M.FNN_IP_L5, ("ip", 7): M.FNN_IP_L7}[(cross, len(hidden))]
    if F == 39:
        cat_sizes = data.CRITEO_CAT_SIZES if X == 13 + int(data.CRITEO_CAT_SIZES.sum()) else data.scaled_criteo_cat_sizes(X)
        n_num = 13
    else:
        cat_sizes, n_num = data.scaled_criteo_cat_sizes(X + 13) if F == 26 else data.ipinyou_field_sizes(X), 0
    offsets = data.offsets_of(cat_sizes, n_num)
    m = cls(cat_sizes, offsets, N, [X, F, rank] + hidden + ["tanh"], ["uniform", -0.01, 0.01, list(range(1, 20)), None],
            ["adam", 1e-4, 1e-8, "sum"] if wl.endswith("_adam") else ["sgd", 1e-3, "sum"], [1.0], "train", N, shard=shard)
    for l in range(m.engine.nl):     # +-0.01 weights make every activation ~0: use fan-in scaling for a realistic step
        m.engine.W(l).uniform_(-1, 1).mul_(1.0 / np.sqrt(m.engine.widths[l]))

    def train(ids, wts, y):
        return m.train_step({m.v_wt_hldr: wts[:, :n_num], m.c_id_hldr: ids[:, n_num:] - offsets, m.c_wt_hldr: wts[:, n_num:], m.lbl_hldr: y})

    def feed_of(ids, wts, y):        # the reference's feed: numeric weights, per-field LOCAL categorical ids, weights, labels
        return {m.v_wt_hldr: np.ascontiguousarray(wts[:, :n_num]), m.c_id_hldr: np.ascontiguousarray(ids[:, n_num:] - offsets),
                m.c_wt_hldr: np.ascontiguousarray(wts[:, n_num:]), m.lbl_hldr: y}
    train.feed_of, train.step = feed_of, m.train_step
    return m, train


# algorithmic bytes / flops per call; args are the ctypes arguments of the call
def call_cost(name, a):
    if name in ("dctr_gemm", "dctr_gemm_tc"):
        return "flop", 2.0 * a[1] * a[2] * a[3]
    if name == "dctr_gather_concat_fwd":
        N, F, d, ldz = a[5], a[6], a[7], a[13]
        cross = a[8]
        D0 = F * d + (F * (F - 1) // 2 if cross == 1 else d * d if cross == 2 else 0) + 1
        return "byte", N * (F * 8 + F * 4 * d + 4 * D0)
    if name == "dctr_fm_fwd":
        N, F, d = a[5], a[6], a[7]
        return "byte", N * (F * 8 + 4 + F * 4 * d + 4 + 4 * d)
    if name in ("dctr_scatter_sgd", "dctr_scatter_sgd_masked", "dctr_scatter_sgd_fm"):
        M, d = (a[2], a[4]) if name == "dctr_scatter_sgd_fm" else (a[3], a[5])
        return "byte", M * (4 + 4 * d + 2 * 4 * d)
    if name == "dctr_cross_bwd":
        N, F, d, ldz = a[5], a[6], a[7], a[10]
        return "byte", N * (F * 8 + F * 4 * d + 4 * ldz + F * 4 * d)
    if name in ("dctr_sort_rows", "dctr_sort_pairs"):
        M, X = (a[1], a[2]) if name == "dctr_sort_rows" else (a[2], a[3])
        bits = max(1, int(X - 1).bit_length())
        passes = (bits + 7) // 8
        if M <= (1 << 18):
            passes = min(passes, (bits + 9) // 10)           # short lists: 10-bit digits when they save a pass
        return "byte", M * 4 + passes * M * 16     # one histogram read of the keys + per pass read/write of (key, pos)
    if name == "dctr_splitk_reduce":
        return "byte", (a[1] + 1) * a[2] * a[3] * 4
    if name == "dctr_dense_sgd":
        return "byte", a[2] * 12
    if name == "dctr_out_bwd":
        return "byte", a[4] * a[5] * 8
    if name == "dctr_col_wsum":
        return "byte", a[2] * a[3] * 4
    if name == "dctr_out_xent_fwd":
        return "byte", a[2] * a[3] * 4
    if name == "dctr_vec_sum":
        return "byte", a[1] * 4
    if name == "dctr_out_bwd_fused":                # reads the last activation, writes its gradient (+ dlogit)
        return "byte", a[4] * a[5] * 8 + a[4] * 4
    if name == "dctr_arrange_fields":               # feats in, ids + wts out
        return "byte", a[2] * (a[3] * 4 + a[4] * 8)
    if name == "dctr_mlp2tc_train":                 # 3xTF32 on tcgen05: forward + data gradients of both hidden layers (fp32-equivalent flops)
        return "flop", 4.0 * a[0] * (a[1] * a[2] + a[2] * a[3] + a[3])
    if name == "dctr_mlp2tc_pack":                  # W1, W2 read; each written twice (both orientations) as hi + lo planes
        return "byte", (a[0] * a[1] + a[1] * a[2]) * 4 * 5
    if name == "dctr_splitk_reduce_sgd":            # slabs + W read, W written
        return "byte", (a[1] + 2) * a[2] * a[3] * 4
    if name == "dctr_col_wsum_sgd":
        return "byte", a[2] * a[3] * 4
    if name == "dctr_mlp2_train":                   # fp32 FFMA: forward + data gradients of both hidden layers
        return "flop", 4.0 * a[0] * (a[1] * a[2] + a[2] * a[3] + a[3])
    if name == "dctr_mlp2_wgrad":
        return "flop", 2.0 * a[0] * (a[1] * a[2] + a[2] * a[3])
    if name == "dctr_arrange_update_list":          # feats in, ids + wts + update list out
        return "byte", a[2] * (a[3] * 12 + a[4] * 8)
    if name == "dctr_out_unit_train":               # the last activation in, its gradient out, pred / loss / dlogit
        return "byte", a[2] * a[3] * 8 + a[2] * 16
    if name == "dctr_gather_rows":                  # id + row in, row out
        return "byte", a[4] * (4 + 8 * a[3])
    if name == "dctr_dense_adam":                   # p, g, m, v read; p, m, v (+ g) written
        return "byte", a[4] * 4 * (7 + (1 if a[10] else 0))
    return "byte", 0.0


# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed `ncu --set full` captures
# (profiles/r01_ncu_*.txt), keyed by C-ABI call and the workload it was captured on; null for kernels not captured.
NCU_TRAFFIC = {
    ("fnn_ipinyou", "dctr_mlp2tc_train"): 4.677e6 + 768.0,                     # profiles/r02_ncu_mlp2tc_final.txt (weights + x in, L2-resident outputs)
    ("fnn_criteo", "dctr_gemm_tc[NN 65536x400x430]"): 114.06e6 + 70.96e6,      # profiles/r01_ncu_criteo_gemm_tc_kernel_v2.txt
    ("fnn_criteo", "dctr_col_wsum"): 109.07e6 + 4.51e6,                       # profiles/r01_ncu_criteo_col_wsum_kernel.txt
    # hist + 2 wide passes at M = 65536: the lists live in L2, only cold reads reach DRAM (profiles/r01_ncu_fnn_sort_kernels_step_size.txt)
    ("fnn_ipinyou", "dctr_sort_rows"): 285.95e3 + 321.79e3 + 577.28e3,
    ("fm_ipinyou", "dctr_sort_rows"): 285.95e3 + 321.79e3 + 577.28e3,
    ("emb", "gather_concat_fwd"): 257.25e6 + 698.65e6,                        # profiles/r01_ncu_emb_gather_concat_kernel.txt
    ("emb", "fm_fwd"): 234.24e6 + 51.89e6,                                    # profiles/r01_ncu_emb_fm_fwd_kernel.txt
    # profiles/r02_ncu_emb_fold_tile_kernel_final.txt: re-captured on the final fold of round 2 (update list, double-precision decay): the
    # same 4.37 GB the round-1 kernel moved (r01_ncu_emb_fold_tile_kernel_chunk4.txt); without the per-slot weight gather 2.76 GB
    ("emb", "scatter_sgd[sum]"): 4374.34e6 + 42.05e6,
    ("emb", "scatter_sgd_masked[seq_decay]"): 2756.41e6 + 42.46e6,
}


def emb_roofline(N=1 << 20, F=16, rank=10, X=20_000_000, reps=5, zipf=1.1):
    """The HBM-bound kernels alone at a size where they ARE bandwidth-bound (table 0.96 GB >> 126 MB L2,
    N = 2^20 samples per launch): achieved algorithmic GB/s vs the measured copy peak."""
    import torch
    from deepctr_b200 import _lib as L, data
    if os.environ.get("DCTR_L2_FETCH"):       # experiment knob: cudaLimitMaxL2FetchGranularity (enum value 5)
        import ctypes
        rt = ctypes.CDLL("libcudart.so")
        torch.zeros(1, device="cuda")
        val = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(val), 5)
        rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(int(os.environ["DCTR_L2_FETCH"])))
        val2 = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(val2), 5)
        print("L2 fetch granularity: was %d, set rc=%d, now %d" % (val.value, rc, val2.value), file=sys.stderr)
    d = rank + 1
    ld = (d + 3) // 4 * 4
    rng = np.random.default_rng(7)
    sizes = data.ipinyou_field_sizes(X)
    ids_h = data.make_field_batch(rng, N, sizes, zipf).astype(np.int32)        # zipf = 0: uniform within each field
    dev = "cuda"
    ids = torch.as_tensor(ids_h).to(dev)
    wts = torch.ones(N, F, device=dev)
    table = torch.empty(X, ld, device=dev).uniform_(-0.01, 0.01)
    D0 = 1 + F * d
    ldz = (D0 + 3) // 4 * 4
    z = torch.empty(N, ldz, device=dev)
    dz = torch.empty(N, ldz, device=dev).uniform_(-1, 1)
    ge = torch.empty(N, F, ld, device=dev).uniform_(-1, 1)
    y = (torch.rand(N, device=dev) < 0.05).float()
    bias = torch.zeros(1, device=dev)
    M = N * F
    rows, pos = torch.empty(M, dtype=torch.int32, device=dev), torch.empty(M, dtype=torch.int32, device=dev)
    sws = torch.empty(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device=dev)
    fws = torch.empty(L.lib().dctr_scatter_workspace(M, ld), dtype=torch.uint8, device=dev)
    logit, pred, lossn, g = (torch.empty(N, device=dev) for _ in range(4))
    s = torch.empty(N, ld, device=dev)
    p = lambda t: t.data_ptr()
    calls = {
        "gather_concat_fwd": ("dctr_gather_concat_fwd", (p(ids), p(wts), p(table), X, ld, N, F, d, 0, 1, p(bias), 0, p(z), ldz, None)),
        "fm_fwd": ("dctr_fm_fwd", (p(ids), p(wts), p(table), X, ld, N, F, d, p(bias), p(y), 1.0, p(logit), p(pred), p(lossn), p(g), p(s), ld, None)),
        "sort_rows": ("dctr_sort_rows", (p(ids), M, X, p(rows), p(pos), p(sws), sws.numel(), None)),
        "scatter_sgd[seq_decay]": ("dctr_scatter_sgd", (1, p(rows), p(pos), M, F, d, p(dz) + 4, ldz, d, p(wts), 1e-3, 0.9999, p(table), X, ld, p(fws), fws.numel(), None)),
        "scatter_sgd[sum]": ("dctr_scatter_sgd", (0, p(rows), p(pos), M, F, d, p(dz) + 4, ldz, d, p(wts), 1e-3, 1.0, p(table), X, ld, p(fws), fws.numel(), None)),
        # 0 / 1 weights with the 0s on one pad row (FNN_wnzh, SNN): no per-slot weight gather
        "scatter_sgd_masked[seq_decay]": ("dctr_scatter_sgd_masked", (1, p(rows), p(pos), M, F, d, p(dz) + 4, ldz, d, X - 1, 1e-3, 0.9999, p(table), X, ld, p(fws), fws.numel(), None)),
        "scatter_sgd_masked[sum]": ("dctr_scatter_sgd_masked", (0, p(rows), p(pos), M, F, d, p(dz) + 4, ldz, d, X - 1, 1e-3, 1.0, p(table), X, ld, p(fws), fws.numel(), None)),
        # the same fold fed from a per-slot gradient buffer [N, F, ld] (what the IP / OP models and the multi-GPU steps hand it):
        # 16-byte aligned 48-byte rows, 128-bit loads, weights already applied
        "scatter_sgd[sum, slot buffer]": ("dctr_scatter_sgd", (0, p(rows), p(pos), M, F, d, p(ge), F * ld, ld, None, 1e-3, 1.0, p(table), X, ld, p(fws), fws.numel(), None)),
        "scatter_sgd_fm": ("dctr_scatter_sgd_fm", (p(rows), p(pos), M, F, d, p(g), p(s), ld, p(wts), 1e-3, 1.0, p(table), X, ld, p(fws), fws.numel(), None)),
    }
    pk = peaks()
    out = {"N": N, "F": F, "d": d, "rows": X, "ids": "zipf(%g) within each field" % zipf if zipf else "uniform within each field",
           "unique_rows": int(len(np.unique(ids_h))), "table_GB": round(X * ld * 4 / 1e9, 2), "peak_GBs": pk["hbm"], "kernels": {}}
    for key, (name, a) in calls.items():
        L.call(name, *a)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            L.call(name, *a)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        _, cost = call_cost(name, a)
        gbs = cost / (ms * 1e-3) / 1e9
        out["kernels"][key] = {"ms": round(ms, 4), "algorithmic_MB": round(cost / 1e6, 1), "GBs": round(gbs, 1), "frac": round(gbs / pk["hbm"], 4)}
        tr = NCU_TRAFFIC.get(("emb", key))
        if tr is not None and (N, F, rank, X, zipf) == (1 << 20, 16, 10, 20_000_000, 1.1):       # the configuration the captures were taken on
            out["kernels"][key]["ncu_dram_MB"] = round(tr / 1e6, 1)
    return out


def run_b200(args, rank, world):
    import torch
    import deepctr_b200  # noqa: F401
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = measure_workload(args.workload, args.batch, args.steps, max(3, args.warmup), rank, world, dist, local, not args.no_profile)
    if rank == 0 and world == 1 and not args.no_profile and not args.no_emb:
        torch.cuda.empty_cache()
        out["emb_roofline"] = emb_roofline()
        out["emb_roofline_uniform_ids"] = emb_roofline(zipf=0.0)
    # the other configurations BASELINE.json names, at the same number of GPUs, in the same driver run: config 5 (Criteo-shaped,
    # 10^8-row table row-sharded over the ranks: the curve the north star's 0.8 scaling target is stated on) and config 4
    # (FNN_IP_L7 at batch 65 536 per GPU, data parallel)
    if args.workload == "fnn_ipinyou" and not args.batch and not args.no_secondary:
        out["secondary"] = {}
        for wl2 in ("fnn_criteo_sharded", "fnn_ip_l7_ipinyou"):
            r2 = measure_workload(wl2, 0, min(args.steps, 20), 3, rank, world, dist, local, False)
            out["secondary"][wl2] = {k: r2[k] for k in ("value", "unit", "ms_per_step", "n_gpus", "config", "e2e", "e2e_pipelined", "launches_per_step", "clocks")}
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(args.workload, args.batch or WORKLOADS[args.workload][6], budget_s=12.0)
    if rank == 0:
        print_json(out)
    if dist:
        from deepctr_b200 import dist as ddist
        ddist.shutdown(dist, [])


def measure_workload(wl, batch, steps, warmup, rank, world, dist, local, profile):
    """One workload at `world` GPUs: device-timed steps (value), the end-to-end loop through the model class (e2e), the
    per-kernel pass (profile).  Returns the JSON line as a dict."""
    import torch
    from deepctr_b200 import _lib
    import gc
    form, F, X, rank_k, hidden, cross, N = WORKLOADS[wl]
    N = batch or N
    sharded = wl.endswith("_sharded")
    model, train = build_model(wl, N, shard=(rank, world) if sharded else None)
    if world > 1:
        from deepctr_b200 import dist as ddist
        if sharded:
            ddist.attach_sharded(model, dist, X)
        else:
            ddist.attach(model, dist)
    eng = model.engine
    pool, _ = make_pool(form, F, X, N, 8, seed=1234 + rank)
    dpool = [(torch.as_tensor(i.astype(np.int32)).cuda(), torch.as_tensor(w).cuda(), torch.as_tensor(y).cuda()) for i, w, y in pool]
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    K, W = steps, warmup

    def stage_inputs(i):
        """the batch of step i into the step's input buffers: `value` is defined with the inputs already resident in HBM when the timed
        region starts, so this runs BEFORE the L2 flush and outside the timed region (the step then reads its inputs cold)"""
        ids, wts, y = dpool[i % len(dpool)]
        if form == "theano":
            model.feats.copy_(ids)      # raw feature lists; the slot-per-field arrangement is part of the device step
        else:
            eng.ids.copy_(ids)
            eng.wts.copy_(wts)
        eng.y.copy_(y)

    def run_step():
        model._pre_step()               # this step's dropout masks (host generator of the Theano form) ...
        if form == "theano":
            model._upload_masks()       # ... and their upload are part of the step
        model._run_step()

    def device_step(i):
        stage_inputs(i)
        run_step()

    sampler = ClockSampler(local)
    sampler.start()
    for i in range(W):
        device_step(i)
    torch.cuda.synchronize()
    assert model._graph is not None or not model.use_graph
    l0 = _lib.launch_count()
    was_graph, g = model.use_graph, model._graph
    model.use_graph = False
    device_step(0)                       # one eager step just to count this library's launches per step
    per_step_launches = _lib.launch_count() - l0
    model.use_graph, model._graph = was_graph, g
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.begin()
    evs = []
    for i in range(K):
        stage_inputs(W + i)              # inputs resident in HBM ...
        flush.zero_()                    # ... and evicted from L2 with the table / weights / activations (not timed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run_step()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    # warm-L2 variant (no flush, back to back): reported next to the cold number
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        device_step(W + i)
    e1.record()
    torch.cuda.synchronize()
    warm_ms = e0.elapsed_time(e1)
    clocks = sampler.summary()
    shard_phases = None
    if getattr(model, "sharded", None) is not None and os.environ.get("DCTR_SHARD_PROFILE"):
        was_graph, g = model.use_graph, model._graph        # eager steps with CUDA events between the phases of the exchange
        model.use_graph = False
        model.sharded.summary()
        for i in range(5):
            device_step(W + i)
        shard_phases = dict(model.sharded.summary(), capacity=model.sharded.C, slots=eng.M)
        model.use_graph, model._graph = was_graph, g
    if dist:
        t = torch.tensor([total_ms, warm_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, warm_ms = float(t[0]), float(t[1])

    # ---- end to end: host numpy in, loss / preds out, through the model class ----
    if hasattr(train, "feed_of"):      # batches in the form the reference's input pipeline hands to Session.run
        hp = [(train.feed_of(*b),) for b in pool]
        train = train.step
    else:
        hp = pool
    for i in range(3):
        train(*hp[i % len(hp)])
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(K):
        loss, preds = train(*hp[i % len(hp)])
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if dist:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t[0])
    assert np.isfinite(loss) and np.isfinite(preds).all()

    # ---- the same loop through the two-deep pipelined call (train_async / train_step_async: the result of step t comes back from call
    # t + 1; pack + upload of batch t + 1 run under step t).  Every step still uploads its inputs from pinned memory and reads its
    # pred | loss back inside the timed region; an addition next to `e2e`, which stays the reference's synchronous call
    e2e_pipe = None
    try:
        if dist:
            raise NotImplementedError("measured on one GPU only (the multi-GPU steps have not been run through the pipelined call)")
        if form == "theano":
            astep = lambda ids, wts, y: model.train_async(ids, y)
        elif form == "fm":
            astep = lambda ids, wts, y: model.train_step_async({model.sp_id_hldr: ids.reshape(-1), model.sp_wt_hldr: wts.reshape(-1), model.lbl_hldr: y})
        else:
            astep = model.train_step_async
        for i in range(3):
            astep(*hp[i % len(hp)])
        model.flush_async()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(K):
            astep(*hp[i % len(hp)])
        loss_p, preds_p = model.flush_async()
        torch.cuda.synchronize()
        pipe_s = time.perf_counter() - t0
        assert np.isfinite(loss_p) and np.isfinite(preds_p).all()
        e2e_pipe = {"value": N * world * K / pipe_s, "unit": "samples/s", "ms_per_step": pipe_s / K * 1e3,
                    "h2d_bytes_per_step": model.h2d_bytes, "d2h_bytes_per_step": model.d2h_bytes,
                    "call": "train_async / train_step_async + flush_async (two-deep pipeline; results one call late)"}
    except Exception as ex:                                         # noqa: BLE001  (an extra measurement must not cost the line)
        e2e_pipe = {"unavailable": "%s: %s" % (type(ex).__name__, str(ex)[:200])}

    # ---- per-kernel timing (CUDA events on the launching stream; eager; launches queued behind a sleep) ----
    kernels, roof = [], None
    if profile:
        was_graph = model.use_graph
        model.use_graph = False
        eng.overlap_wgrad = False
        eng.overlap_sort = False           # the row sort normally overlaps the MLP on a side stream: serialise it to time it alone
        agg = {}
        P = 5
        for i in range(P):
            flush.zero_()
            torch.cuda._sleep(2_000_000)           # ~1 ms: lets the host enqueue the whole step ahead of the GPU
            _lib.PROFILE = []
            device_step(i)
            torch.cuda.synchronize()
            for name, a, s, e in _lib.PROFILE:
                key = name if not name.startswith("dctr_gemm") else "%s[%s %dx%dx%d]" % (name, ["NN", "NT", "TN"][a[0]], a[1], a[2], a[3])
                unit, cost = call_cost(name, a)
                r = agg.setdefault(key, dict(ms=0.0, cost=0.0, unit=unit, calls=0))
                r["ms"] += s.elapsed_time(e)
                r["cost"] += cost
                r["calls"] += 1
            _lib.PROFILE = None
        model.use_graph = was_graph
        eng.overlap_sort = True
        eng.overlap_wgrad = True
        pk = peaks()
        tot = sum(r["ms"] for r in agg.values())
        for key, r in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
            ach = r["cost"] / (r["ms"] * 1e-3) / (1e12 if r["unit"] == "flop" else 1e9) if r["ms"] > 0 else 0.0
            peak = pk["tensor"] if r["unit"] == "flop" else pk["hbm"]
            kernels.append({"kernel": key, "us_per_step": round(r["ms"] / P * 1e3, 2), "share": round(r["ms"] / tot, 3),
                            "achieved": round(ach, 2), "unit": "TFLOP/s" if r["unit"] == "flop" else "GB/s",
                            "frac": round(ach / peak, 4), "calls_per_step": r["calls"] // P})
        top = kernels[0]
        roof = {"kernel": top["kernel"], "bound": "tensor" if top["unit"] == "TFLOP/s" else "hbm", "achieved": top["achieved"],
                "peak": pk["tensor"] if top["unit"] == "TFLOP/s" else pk["hbm"], "unit": top["unit"], "frac": top["frac"],
                "traffic": NCU_TRAFFIC.get((wl, top["kernel"])), "peak_source": pk["src"] + (" (bf16 sustained)" if top["unit"] == "TFLOP/s" else " (copy)"),
                "share_of_step": top["share"]}
        if top["unit"] == "TFLOP/s" and top["kernel"].startswith("dctr_gemm_tc"):
            # achieved counts fp32-equivalent flops (2 M N K); the kernel issues three tf32 MMAs per k-step (error compensation)
            roof["note"] = ("3xTF32 on tcgen05: fp32-equivalent FLOP/s; its ceiling is cuBLAS TF32 / 3 = 243 TFLOP/s (DESIGN 3.4) "
                            "-> frac_of_3xtf32_ceiling; a GEMM of this size is launch- and latency-bound")
            roof["frac_of_3xtf32_ceiling"] = round(top["achieved"] / 243.0, 4)
        elif top["unit"] == "TFLOP/s" and top["kernel"].startswith("dctr_mlp2tc"):
            # 64 CTAs of 64 samples issue M = 128 MMAs over half-empty tiles on 64 of the 148 SMs: the ceiling that applies to THIS launch
            # is 243 TFLOP/s x 64 / 148 x 1 / 2 = 52.5 TFLOP/s fp32-equivalent; the epilogues (activation, tf32 split, the transposition
            # through shared memory: ~25 instructions per element) bound it, not the MMAs
            roof["note"] = ("forward + data gradients of both hidden layers chained through tcgen05 in one kernel (csrc/mlp2tc.cu), 3xTF32: "
                            "fp32-equivalent FLOP/s; ceiling of 3xTF32 = 243 TFLOP/s -> frac_of_3xtf32_ceiling; epilogue-bound (DESIGN 3.5)")
            roof["frac_of_3xtf32_ceiling"] = round(top["achieved"] / 243.0, 4)
        elif top["unit"] == "TFLOP/s" and top["kernel"].startswith("dctr_mlp2"):
            roof["note"] = ("fused fp32 FFMA kernel (csrc/mlp2.cu); the layers of this net are too small for tensor-core tiles, so the "
                            "bound that applies is the fp32 SIMT peak, 148 SMs x 128 lanes x 2 x 1.965 GHz = 74.4 TFLOP/s -> frac_of_fp32_peak")
            roof["frac_of_fp32_peak"] = round(top["achieved"] / 74.4, 4)
        elif top["unit"] == "TFLOP/s":
            roof["note"] = "fp32 FFMA parity path; fp32 SIMT peak is 74.4 TFLOP/s -> frac_of_fp32_peak"
            roof["frac_of_fp32_peak"] = round(top["achieved"] / 74.4, 4)

    out = {
        "metric": "FNN train samples/s", "value": N * world * K / (total_ms * 1e-3), "unit": "samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "form": form, "fields": F, "rows": X, "rank": rank_k, "hidden": hidden, "cross": cross,
                   "batch_per_gpu": N, "global_batch": N * world, "l2": "flushed (512 MB write) before every timed step",
                   "parallelism": ("rowshard%d" if sharded else "dp%d") % world if world > 1 else "single", "cuda_graph": bool(model.use_graph)},
        "warm_l2": {"ms_per_step": warm_ms / K, "value": N * world * K / (warm_ms * 1e-3)},
        **({"shard_phases_ms": shard_phases} if shard_phases else {}),
        "e2e": {"value": N * world * K / e2e_s, "unit": "samples/s", "h2d_bytes_per_step": model.h2d_bytes,
                "d2h_bytes_per_step": model.d2h_bytes, "ms_per_step": e2e_s / K * 1e3},
        "e2e_pipelined": e2e_pipe,
        "gpu_launches": int(per_step_launches * K), "launches_per_step": int(per_step_launches),
        "clocks": clocks,
    }
    if roof:
        out["roofline"] = roof
        out["kernels"] = kernels
    # release the model (and its captured graph) before the next workload is built
    model._graph = None
    del flush, dpool, model, eng, train
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement of the reference step (numpy float64 + OpenBLAS, all host cores)
# ---------------------------------------------------------------------------------------------------
def _oracle_on_path():
    """oracle/ is test infrastructure: only the CPU legs of this file (cpu_baseline, --impl reference) put it on the path"""
    p = os.path.join(ROOT, "oracle")
    if p not in sys.path:
        sys.path.insert(0, p)


def oracle_stepper(wl, N, seed=1234, dtype=np.float64):
    _oracle_on_path()
    import deepctr_oracle as O
    form, F, X, rank, hidden, cross, _ = WORKLOADS[wl]
    rng = np.random.default_rng(seed)
    d = rank + 1
    table = rng.uniform(-0.01, 0.01, (X, d))
    if form == "fm":
        fm = O.OracleFM(table, 0.0, dtype=dtype)
        return lambda ids, wts, y: fm.train_step(ids, wts, y, 1e-3, 0.0, dense_l2=False)
    cw = {"none": 0, "ip": F * (F - 1) // 2, "op": d * d}[cross]
    dims = [F * d + cw + 1] + hidden + [1]
    Ws = [rng.uniform(-1, 1, (dims[i], dims[i + 1])) / np.sqrt(dims[i]) for i in range(len(dims) - 1)]
    bs = [np.zeros(dims[i + 1]) for i in range(len(dims) - 1)]
    if form == "theano":
        net = O.OracleNet(table, 0.0, Ws, bs, F, d, "none", True, "none", "tanh", dtype=dtype)
        mrng = np.random.RandomState(234)

        def step(ids, wts, y):
            masks = [None, mrng.binomial(1, 0.5, (1, hidden[0])).astype(dtype), mrng.binomial(1, 0.5, (1, hidden[1])).astype(dtype)]
            return net.train_step_theano(ids, y, 1e-3, 0.1, masks=masks, closed_form=True)
        return step
    net = O.OracleNet(table, 0.0, Ws, bs, F, d, cross, False, "tanh", "tanh", dtype=dtype)
    if wl.endswith("_adam"):
        return lambda ids, wts, y: net.train_step_tf_adam(ids, wts, y, 1e-4, 1e-8)
    return lambda ids, wts, y: net.train_step_tf(ids, wts, y, 1e-3)


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to its workers: give the CPU arm's BLAS every host core back."""
    n = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=n)
    except Exception:
        pass
    return n


def faithful_baseline(wl, n1=512, budget_s=4.0):
    """B1 of SURVEY 8(d): what the reference COSTS -- its own structure on a reduced batch: the per-sample, per-feature gather of
    feats_to_layer_one_array (FNN_wnzh.py:87-96), BLAS for the Theano graph, and the interpreted triple update loop (:298-306)."""
    _oracle_on_path()
    import deepctr_oracle as O
    form, F, X, rank, hidden, cross, _ = WORKLOADS[wl]
    if form != "theano":
        return None
    from deepctr_b200 import data
    rng = np.random.default_rng(5)
    d = rank + 1
    sizes = data.ipinyou_field_sizes(X)
    feat_field = np.repeat(np.arange(F), sizes)
    table = np.concatenate([rng.uniform(-0.01, 0.01, (X, d)), np.zeros((1, d))], 0)
    dims = [F * d + 1] + hidden + [1]
    Ws = [rng.uniform(-1, 1, (dims[i], dims[i + 1])) / np.sqrt(dims[i]) for i in range(3)]
    net = O.OracleNet(table, 0.0, Ws, [np.zeros(dims[i + 1]) for i in range(3)], F, d, "none", True, "none", "tanh")
    n, t0 = 0, time.perf_counter()
    while True:
        ids = data.make_field_batch(rng, n1, sizes, 1.1)
        y = (rng.random(n1) < 0.05).astype(np.float64)
        feats = [list(r) for r in ids]
        x = np.zeros((n1, dims[0]))
        for t, ft in enumerate(feats):                       # feats_to_layer_one_array, one sample at a time
            x[t, 0] = 0.0
            for feat in ft:
                o = 1 + feat_field[feat] * d
                x[t, o:o + d] = net.table[feat]
        masks = [None, rng.binomial(1, 0.5, (1, hidden[0])).astype(np.float64), rng.binomial(1, 0.5, (1, hidden[1])).astype(np.float64)]
        net.train_step_theano_feats(feats, feat_field, y, 1e-3, 0.1, masks)       # BLAS graph + the literal triple loop
        n += 1
        el = time.perf_counter() - t0
        if el > budget_s:
            break
    return {"value": n * n1 / el, "unit": "samples/s", "kind": "port", "cores": 1,
            "sample": "%d steps of batch %d: per-sample gather loop + BLAS graph + interpreted triple update loop (the reference's own structure)" % (n, n1)}


def reference_script_baseline(budget_s=8.0, n_train=2000, n_test=500):
    """reference_script_run() guarded: this leg is extra evidence and must never cost the bench its JSON line."""
    try:
        return reference_script_run(budget_s, n_train, n_test)
    except Exception as ex:                                         # noqa: BLE001
        return {"unavailable": "%s: %s" % (type(ex).__name__, str(ex)[:200]), "kind": "reference"}


def reference_script_run(budget_s=8.0, n_train=2000, n_test=500):
    """B0: the reference's OWN script timed on the host -- python/FNN_wnzh.py as converted by oracle/build_ref.py (mechanical Py2->3
    only), executed top to bottom on the Theano stand-in of oracle/shims (torch CPU float64 for the compiled graph), with its
    hard-coded advertiser-2997 hyper-parameters (batch 100, FNN_wnzh.py:18,43-49) on the S1 demo set of SURVEY 8(d) written in its
    text formats.  One loop iteration = get_batch_data (linecache + split + dict gather, :224-253), train() (:177-182) and the
    interpreted triple update loop (:298-306); timed from the exit of one train() call to the exit of the next inside an epoch, so
    the per-epoch evaluation passes are not counted.  Returns None when oracle/_ref is not there."""
    _oracle_on_path()
    import tempfile
    import ref_run
    if not ref_run.available():
        return None
    from deepctr_b200 import data
    rng = np.random.default_rng(1234)
    F, d, X = 16, 11, 2048
    sizes = data.ipinyou_field_sizes(X)
    feat_field = np.repeat(np.arange(F), sizes)
    table = rng.uniform(-0.05, 0.05, (X, d))
    wd = tempfile.mkdtemp(prefix="dctr_ref_script_")
    os.makedirs(os.path.join(wd, "data"))
    data.write_fm_model(os.path.join(wd, "data", "fm.model.txt"), 0.1, table, feat_field)
    for name, n in (("train", n_train), ("test", n_test)):
        ids = data.make_field_batch(rng, n, sizes)
        y = data.planted_labels(rng, ids, table * 20, pos_rate=0.2)
        with open(os.path.join(wd, "data", "%s.fm.txt" % name), "w") as f:
            for t in range(n):
                f.write("%d %s\n" % (y[t], " ".join("%d:1" % i for i in ids[t])))
    try:
        _, calls = ref_run.run_fnn_wnzh(wd, budget_s=budget_s)
    finally:
        import shutil
        shutil.rmtree(wd, ignore_errors=True)
    el, n, graph = 0.0, 0, 0.0
    for a, b in zip(calls[:-1], calls[1:]):
        if a["fn"] == "function1" and b["fn"] == "function1":      # consecutive train() calls of one epoch
            el += b["t1"] - a["t1"]
            graph += b["t1"] - b["t0"]                             # the compiled-graph call alone (the stand-in's share of the iteration)
            n += 1
    if not n:
        return None
    N = calls[0]["inputs"][0].shape[0]
    return {"value": n * N / el, "unit": "samples/s", "kind": "reference", "cores": 1, "ms_per_step": el / n * 1e3,
            # the stand-in evaluates the graph symbolically (slower than compiled Theano): the same iterations with the train() call
            # itself taken out = an upper bound on what the script can reach with a free graph engine
            "train_call_ms": graph / n * 1e3, "value_without_train_call": n * N / (el - graph),
            "sample": "%d loop iterations of python/FNN_wnzh.py itself (oracle/_ref: Py2->3 conversion only; Theano stand-in on torch CPU "
                      "float64), batch %d as the script hard-codes, S1 demo set (16 fields, %d features): batch read + dict gather + "
                      "train() + the interpreted triple update loop" % (n, N, X)}


def cpu_baseline(wl, N, budget_s=12.0, steps=None):
    form, F, X, *_ = WORKLOADS[wl]
    use_all_host_threads()
    step = oracle_stepper(wl, N)
    pool, _ = make_pool(form, F, X, N, 4, seed=99)
    step(*pool[0])                                  # warm-up (BLAS threads, page faults)
    n, t0 = 0, time.perf_counter()
    while True:
        step(*pool[n % len(pool)])
        n += 1
        el = time.perf_counter() - t0
        if (steps and n >= steps) or (not steps and el > budget_s):
            break
    out = {"value": n * N / el, "unit": "samples/s", "cores": os.cpu_count(), "kind": "port",
           "sample": "%d steps of batch %d of workload %s (B2: vectorised numpy float64 restatement, OpenBLAS threads = all cores; "
                     "the Theano path of the reference is float64, its TF path float32)" % (n, N, wl),
           "ms_per_step": el / n * 1e3}
    if not steps:
        # the same port in float32 (what the TF generation of the reference computes in) and the faithful loop structure (B1)
        step32 = oracle_stepper(wl, N, dtype=np.float32)
        step32(*pool[0])
        k, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < budget_s / 3:
            step32(*pool[k % len(pool)])
            k += 1
        out["float32"] = {"value": k * N / (time.perf_counter() - t0), "unit": "samples/s", "steps": k}
        b1 = faithful_baseline(wl)
        if b1:
            out["faithful_B1"] = b1
        b0 = reference_script_baseline(budget_s=budget_s / 2) if form == "theano" else None
        if b0:
            out["reference_script_B0"] = b0
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path on rank 0's host cores.  The headline value is the vectorised float64 port of the
    step (oracle/deepctr_oracle.py, pinned against the reference's own code by tests/test_oracle_vs_ref.py) on THIS arm's config with
    every host core -- the strongest CPU baseline; the reference's own script (oracle/_ref on the Theano stand-in, its hard-coded
    batch of 100) is timed beside it as cpu_baseline.reference_script_B0 and is ~10x slower than the port."""
    if rank != 0:
        return
    wl = args.workload
    N = args.batch or WORKLOADS[wl][6]
    K, W = args.steps, max(1, args.warmup)
    cores = use_all_host_threads()
    step = oracle_stepper(wl, N)
    form, F, X, rank_k, hidden, cross, _ = WORKLOADS[wl]
    pool, _ = make_pool(form, F, X, N, 4, seed=99)
    for i in range(min(W, 3)):
        step(*pool[i % 4])
    K = min(K, 200)
    t0 = time.perf_counter()
    done = 0
    for i in range(K):
        step(*pool[i % 4])
        done += 1
        if time.perf_counter() - t0 > 120:        # bounded sample
            break
    el = time.perf_counter() - t0
    v = done * N / el
    b0 = reference_script_baseline(budget_s=8.0) if form == "theano" else None
    print_json({
        "impl": "reference", "metric": "FNN train samples/s", "value": v, "unit": "samples/s", "n_gpus": world, "steps": done,
        "warmup": W, "ms_per_step": el / done * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl, "form": form, "fields": F, "rows": X, "rank": rank_k, "hidden": hidden, "cross": cross,
                   "batch_per_gpu": N, "global_batch": N,
                   "note": "vectorised CPU port of the reference step on rank 0's host cores at this config (the stronger baseline); the "
                           "reference's own Py2 / Theano script, converted and run on a Theano stand-in at its hard-coded batch of 100, is "
                           "cpu_baseline.reference_script_B0"},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
                         "sample": "%d steps of batch %d (numpy float64 + OpenBLAS)" % (done, N),
                         **({"reference_script_B0": b0} if b0 else {})},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


print_json = None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="fnn_ipinyou", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-emb", action="store_true", help="skip the embedding-kernel bandwidth sweeps")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary block (BASELINE configs 4 and 5 at the same GPU count)")
    ap.add_argument("--emb-only", action="store_true", help="only run the embedding-kernel roofline sweep")
    ap.add_argument("--emb-n", type=int, default=1 << 20)
    ap.add_argument("--emb-reps", type=int, default=5)
    ap.add_argument("--emb-zipf", type=float, default=1.1, help="Zipf exponent of the ids inside each field; 0 = uniform")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    # stdout carries ONE JSON line: anything libraries print there (NCCL's version banner, build messages) goes to stderr
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    global print_json
    print_json = lambda obj: (real_stdout.write(json.dumps(obj) + "\n"), real_stdout.flush())
    if args.emb_only:
        print_json({"emb_roofline": emb_roofline(N=args.emb_n, reps=args.emb_reps, zipf=args.emb_zipf)})
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world)


if __name__ == "__main__":
    main()
