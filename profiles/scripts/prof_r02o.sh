# round 2, GPU call O: the tcgen05 fused two-hidden-layer kernel, second design (coalesced epilogues through the A slots)
set -x
cd /root/repo
timeout -s KILL 300 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu -x > gpurun_out/r02_mlp2tc_tests_o.txt 2>&1; echo pytest rc=$?
tail -30 gpurun_out/r02_mlp2tc_tests_o.txt | cut -c1-300
timeout -s KILL 120 python tests/m2tc_time.py 4096 2>&1 | tail -2
DCTR_M2TC_TIMING=1 timeout -s KILL 120 python -m pytest tests/test_mlp2tc_gpu.py -q -m gpu -k "deterministic" -s 2>&1 | grep "timing" | tee gpurun_out/r02_mlp2tc_timeline_o.txt
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm" 2>&1 | tail -2
DCTR_MLP2=tc timeout -s KILL 300 python bench.py --steps 200 --warmup 10 --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_o_mlp2tc.json 2> gpurun_out/r02_bench_o_mlp2tc.err; echo bench rc=$?
tail -3 gpurun_out/r02_bench_o_mlp2tc.err | cut -c1-300
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r02_bench_o_mlp2tc.json"))
    print(d["ms_per_step"], d["value"], d["e2e"], d["launches_per_step"])
    for k in d["kernels"]: print("  ", k["kernel"], k["us_per_step"])
except Exception as e:
    print("no bench line", e)
PY
timeout -s KILL 600 python -m pytest tests/test_models_gpu.py tests/test_ref_golden_gpu.py -q -m gpu -k "mlp2tc" > gpurun_out/r02_model_tests_o.txt 2>&1; echo model pytest rc=$?
tail -30 gpurun_out/r02_model_tests_o.txt | cut -c1-300
