set -x
cd /root/repo
timeout -s KILL 300 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu 2>&1 | tail -8 | cut -c1-250
DCTR_M2TC_TIMING=1 timeout -s KILL 120 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu -k "deterministic" -s 2>&1 | grep "timing" | tail -1
timeout -s KILL 600 python -m pytest tests/test_models_gpu.py tests/test_ref_golden_gpu.py tests/test_kernels_gpu.py -q -m gpu -k "mlp2tc or gemm or col_wsum or reduc" 2>&1 | tail -12 | cut -c1-250
for mode in tc; do
DCTR_MLP2=$mode timeout -s KILL 300 python bench.py --steps 200 --warmup 10 --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_p_$mode.json 2> gpurun_out/r02_bench_p_$mode.err; echo bench rc=$?
tail -3 gpurun_out/r02_bench_p_$mode.err | cut -c1-300
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r02_bench_p_$mode.json"))
    print("$mode", d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["launches_per_step"], d.get("warm_l2"))
    for k in d["kernels"]: print("  ", k["kernel"], k["us_per_step"])
except Exception as e:
    print("no bench line", e)
PY
done
