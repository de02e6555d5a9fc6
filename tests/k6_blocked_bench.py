"""K6 (sort + fold) at the embedding-sweep size, as ONE global list and L2-BLOCKED: the batch cut into B blocks of N/B samples whose
gradient buffer fits the 126 MB L2; every block is sorted on its own (side streams, independent workspaces) and folded in block
order on the main stream -- so the random 44-byte gradient reads of a block hit lines its other slots already brought in
(profiles/r02_ncu_emb_fold_tile_kernel_final.txt: the global fold moves 2.8-4.4 GB for 1 GB of useful bytes, one 128-byte line per
random access).  Block order = sample order, so SUM and SEQ_DECAY keep the global left-fold semantics.
Run on the GPU box:  python tests/k6_blocked_bench.py [reps]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepctr_b200 import _lib as L, data  # noqa: E402


def main(reps=5, N=1 << 20, F=16, rank=10, X=20_000_000, zipf=1.1):
    d = rank + 1
    ld = (d + 3) // 4 * 4
    rng = np.random.default_rng(7)
    sizes = data.ipinyou_field_sizes(X)
    ids_h = data.make_field_batch(rng, N, sizes, zipf).astype(np.int32)
    dev = "cuda"
    ids = torch.as_tensor(ids_h).to(dev)
    table0 = torch.empty(X, ld, device=dev).uniform_(-0.01, 0.01)
    table0[:, d:] = 0
    table = table0.clone()
    D0 = 1 + F * d
    ldz = (D0 + 3) // 4 * 4
    dz = torch.empty(N, ldz, device=dev).uniform_(-1, 1)
    M = N * F
    p = lambda t: t.data_ptr()
    lib = L.lib()
    main_s = torch.cuda.current_stream()
    alg_bytes = N * F * (4 + 4 + 4 * d + 8 * d) + N * F * (4 + 3 * 16)      # fold + sort (4 + passes x 16, 3 passes here)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    out = {"N": N, "F": F, "rows": X, "algorithmic_MB": round(alg_bytes / 1e6, 1), "peak_GBs": peak, "variants": {}}

    def run(B, mode, n_streams=4):
        Nb, Mb = N // B, M // B
        rows = [torch.empty(Mb, dtype=torch.int32, device=dev) for _ in range(B)]
        pos = [torch.empty(Mb, dtype=torch.int32, device=dev) for _ in range(B)]
        sws = [torch.empty(lib.dctr_sort_rows_workspace(Mb), dtype=torch.uint8, device=dev) for _ in range(B)]
        fws = torch.empty(lib.dctr_scatter_workspace(Mb, ld), dtype=torch.uint8, device=dev)
        side = [torch.cuda.Stream() for _ in range(min(n_streams, B))] if B > 1 else [main_s]
        done = [torch.cuda.Event() for _ in range(B)]

        def once(sorts=True, folds=True):
            for s in side:
                s.wait_stream(main_s)
            for b in range(B):
                s = side[b % len(side)]
                if sorts:
                    L.call("dctr_sort_rows", p(ids) + b * Mb * 4, Mb, X, p(rows[b]), p(pos[b]), p(sws[b]), sws[b].numel(), s.cuda_stream)
                done[b].record(s)
            for b in range(B):
                main_s.wait_event(done[b])
                if folds:
                    L.call("dctr_scatter_sgd_masked", mode, p(rows[b]), p(pos[b]), Mb, F, d, p(dz) + 4 + b * Nb * ldz * 4, ldz, d, X - 1,
                           1e-3, 0.9999 if mode else 1.0, p(table), X, ld, p(fws), fws.numel(), main_s.cuda_stream)

        def timed(**kw):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                once(**kw)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps
        table.copy_(table0)
        once()
        torch.cuda.synchronize()
        result = table.clone()
        once()
        ms = timed()
        return ms, result, timed(folds=False), timed(sorts=False)

    for mode, mname in ((0, "sum"), (1, "seq_decay")):
        ref = None
        for B in (1, 4, 8, 16):
            ms, res, ms_sort, ms_fold = run(B, mode)
            if ref is None:
                ref = res
                err = 0.0
            else:       # same left fold, cut at block edges: float32 rounding differs in the partial sums only
                err = float(((res - ref).abs().max() / ref.abs().max()).item())
            gbs = alg_bytes / (ms * 1e-3) / 1e9
            out["variants"]["masked[%s] B=%d" % (mname, B)] = {"ms": round(ms, 4), "GBs": round(gbs, 1), "frac": round(gbs / peak, 4),
                                                                "max_rel_diff_vs_global": err, "sorts_only_ms": round(ms_sort, 4), "folds_only_ms": round(ms_fold, 4)}
            print("masked[%s] blocks=%2d  sort+fold %.4f ms (sorts alone %.4f, folds alone %.4f)  %.0f GB/s  frac %.3f  diff vs global %.2e"
                  % (mname, B, ms, ms_sort, ms_fold, gbs, gbs / peak, err),
                  file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 5)
