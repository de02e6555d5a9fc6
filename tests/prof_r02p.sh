set -x
cd /root/repo
timeout -s KILL 300 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu 2>&1 | tail -3 | cut -c1-200
DCTR_M2TC_TIMING=1 timeout -s KILL 120 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu -k "deterministic" -s 2>&1 | grep "timing" | tail -1
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm" 2>&1 | tail -2
for mode in off tc; do
DCTR_MLP2=$mode timeout -s KILL 300 python bench.py --steps 200 --warmup 10 --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_p_$mode.json 2> gpurun_out/r02_bench_p_$mode.err; echo bench rc=$?
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r02_bench_p_$mode.json"))
    print("$mode", d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["launches_per_step"])
    for k in d["kernels"]: print("  ", k["kernel"], k["us_per_step"])
except Exception as e:
    print("no bench line", e)
PY
done
