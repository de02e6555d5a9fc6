# round 2, GPU call AP: default bench line with the secondary block on the final code (e2e_pipelined in every block); no CPU legs, no
# embedding sweeps (GPU-minutes nearly spent)
set -x
cd /root/repo
( time timeout -k 5 -s KILL 75 python bench.py --steps 50 --no-cpu --no-emb ) > gpurun_out/r02_bench_final_pipe.json 2> gpurun_out/r02_bench_final_pipe.err; echo bench rc=$?
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_final_pipe.json"))
print(d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e_pipelined"], d["gpu_launches"])
for k, v in d["secondary"].items():
    print(k, round(v["ms_per_step"], 3), round(v["e2e"]["ms_per_step"], 3), v.get("e2e_pipelined"))
PY
tail -4 gpurun_out/r02_bench_final_pipe.err | cut -c1-300
