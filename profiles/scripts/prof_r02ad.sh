# round 2, GPU call AD: final 1-GPU records: whole GPU suite, smoke, default bench line, reference arm, launch list of the default step
set -x
cd /root/repo
timeout -k 5 -s KILL 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_gpu_tests_final.txt 2>&1; echo full rc=$?
tail -4 gpurun_out/r02_gpu_tests_final.txt | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
( time timeout -k 5 -s KILL 900 python bench.py ) > gpurun_out/r02_bench_default_final.json 2> gpurun_out/r02_bench_default_final.err; echo bench rc=$?
tail -4 gpurun_out/r02_bench_default_final.err | cut -c1-100
timeout -k 5 -s KILL 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_reference_final.json 2> /dev/null; echo ref rc=$?
for wl in fnn_criteo fm_ipinyou fnn_ip_l3_ipinyou fnn_op_l3_ipinyou; do
timeout -k 5 -s KILL 300 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_final_$wl.json 2> /dev/null; echo $wl rc=$?
done
timeout -k 5 -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_final.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary --no-emb --no-profile > gpurun_out/r02_ncu_ad.log 2>&1; echo ncu rc=$?
python - <<'PY'
import json, glob
d = json.load(open("gpurun_out/r02_bench_default_final.json"))
print(d["ms_per_step"], d["value"], d["warm_l2"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"].get("frac_of_3xtf32_ceiling"))
print({k: (round(v["ms_per_step"], 3), round(v["value"] / 1e6, 2), round(v["e2e"]["ms_per_step"], 2)) for k, v in d["secondary"].items()})
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["float32"]["value"], d["cpu_baseline"]["faithful_B1"]["value"])
for f in sorted(glob.glob("gpurun_out/r02_bench_final_*.json")):
    x = json.load(open(f)); print(f.split("final_")[1], round(x["ms_per_step"], 4), round(x["value"] / 1e6, 2), round(x["warm_l2"]["ms_per_step"], 4), round(x["e2e"]["ms_per_step"], 4), x["launches_per_step"])
print(json.load(open("gpurun_out/r02_bench_reference_final.json"))["value"])
PY
