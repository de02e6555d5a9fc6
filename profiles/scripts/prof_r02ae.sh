set -x
cd /root/repo
for wl in fnn_ipinyou fm_ipinyou; do
timeout -k 5 -s KILL 300 python bench.py --workload $wl --steps 200 --warmup 10 --no-cpu --no-secondary --no-emb --no-profile > gpurun_out/r02_bench_ae_$wl.json 2> /dev/null; echo rc=$?
python - <<PY
import json
d = json.load(open("gpurun_out/r02_bench_ae_$wl.json"))
print("$wl", d["ms_per_step"], d["value"], d["warm_l2"]["ms_per_step"], d["e2e"]["ms_per_step"])
PY
done
