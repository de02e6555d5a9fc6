set -x
cd /root/repo
timeout -k 5 -s KILL 400 python -m pytest tests/test_kernels_gpu.py tests/test_models_gpu.py -q -m gpu -k "gemm or tc" 2>&1 | tail -3 | cut -c1-250
timeout -k 5 -s KILL 200 python tests/tc_check.py bench 2>&1 | tail -8 | cut -c1-200
for wl in fnn_criteo fnn_ip_l7_ipinyou fnn_ipinyou; do
timeout -k 5 -s KILL 300 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu --no-secondary --no-emb --no-profile > gpurun_out/r02_bench_ac_$wl.json 2> /dev/null; echo rc=$?
python - <<PY
import json
d = json.load(open("gpurun_out/r02_bench_ac_$wl.json"))
print("$wl", d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"])
PY
done
