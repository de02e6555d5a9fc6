# round 2, GPU call AF: bench lines after the input staging moved out of the timed region (inputs resident in HBM before the flush)
set -x
cd /root/repo
( time timeout -k 5 -s KILL 900 python bench.py ) > gpurun_out/r02_bench_default_final.json 2> gpurun_out/r02_bench_default_final.err; echo bench rc=$?
for wl in fnn_criteo fm_ipinyou fnn_ip_l3_ipinyou fnn_op_l3_ipinyou; do
timeout -k 5 -s KILL 300 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_final_$wl.json 2> /dev/null; echo $wl rc=$?
done
DCTR_MLP2=off timeout -k 5 -s KILL 300 python bench.py --no-cpu --no-secondary --no-emb > gpurun_out/r02_bench_final_chain.json 2> /dev/null; echo chain rc=$?
python - <<'PY'
import json, glob
d = json.load(open("gpurun_out/r02_bench_default_final.json"))
print(d["ms_per_step"], d["value"], d["warm_l2"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"].get("frac_of_3xtf32_ceiling"))
print({k: (round(v["ms_per_step"], 3), round(v["value"] / 1e6, 2), round(v["e2e"]["ms_per_step"], 2)) for k, v in d["secondary"].items()})
for f in sorted(glob.glob("gpurun_out/r02_bench_final_*.json")):
    x = json.load(open(f)); print(f.split("final_")[1], round(x["ms_per_step"], 4), round(x["value"] / 1e6, 2), round(x["warm_l2"]["ms_per_step"], 4), round(x["e2e"]["ms_per_step"], 4), x["launches_per_step"])
PY
