# round 2, GPU call AK: the whole GPU suite on the final code, then the default bench line
set -x
cd /root/repo
timeout -k 5 -s KILL 560 python -m pytest tests -q -m gpu -x > gpurun_out/r02_gpu_tests_final.txt 2>&1; echo pytest rc=$?
tail -3 gpurun_out/r02_gpu_tests_final.txt | cut -c1-300
( time timeout -k 5 -s KILL 200 python bench.py ) > gpurun_out/r02_bench_default_final2.json 2> gpurun_out/r02_bench_default_final2.err; echo bench rc=$?
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_default_final2.json"))
print(d["ms_per_step"], d["value"], d["warm_l2"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"])
print({k: (round(v["ms_per_step"], 3), round(v["value"] / 1e6, 2)) for k, v in d["secondary"].items()})
PY
