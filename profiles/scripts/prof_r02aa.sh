# round 2, GPU call AA: whole GPU suite and the default bench line at the final commit
set -x
cd /root/repo
timeout -k 5 -s KILL 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_gpu_tests_final.txt 2>&1; echo full rc=$?
tail -8 gpurun_out/r02_gpu_tests_final.txt | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout -k 5 -s KILL 900 python bench.py ) > gpurun_out/r02_bench_default_final.json 2> gpurun_out/r02_bench_default_final.err; echo bench rc=$?
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_default_final.json"))
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"]["traffic"])
print({k: (v["ms_per_step"], v["value"]) for k, v in d["secondary"].items()})
PY
