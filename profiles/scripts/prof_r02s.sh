# round 2, GPU call S: default bench line (fused tcgen05 step) with all blocks; A/B lines; launch list; ncu of the fused kernel
set -x
cd /root/repo
( time timeout -s KILL 900 python bench.py ) > gpurun_out/r02_bench_default_s.json 2> gpurun_out/r02_bench_default_s.err; echo bench rc=$?
tail -4 gpurun_out/r02_bench_default_s.err | cut -c1-200
DCTR_MLP2=off timeout -s KILL 300 python bench.py --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_s_chain.json 2> /dev/null; echo chain rc=$?
timeout -s KILL 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_s_reference.json 2> /dev/null; echo ref rc=$?
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_s.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-emb --no-profile > gpurun_out/r02_ncu_s.log 2>&1; echo ncu rc=$?
python - <<'PY'
import json
for f in ("gpurun_out/r02_bench_default_s.json", "gpurun_out/r02_bench_s_chain.json", "gpurun_out/r02_bench_s_reference.json"):
    try:
        d = json.load(open(f))
        print(f, d.get("ms_per_step"), d.get("value"), d.get("e2e", {}).get("value"), d.get("gpu_launches"), d.get("roofline", {}).get("kernel"), d.get("roofline", {}).get("frac"))
    except Exception as e:
        print(f, "unreadable", e)
PY
