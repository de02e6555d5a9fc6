# round 2, GPU call AJ (8 GPUs): the default bench line with its secondary block, as the driver's scaling run launches it
set -x
cd /root/repo
timeout -k 5 -s KILL 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29580 bench.py --gpus 8 --steps 50 --warmup 5 > gpurun_out/r02_bench_aj_8gpu.json 2> gpurun_out/r02_bench_aj_8gpu.err; echo bench rc=$?
grep -v "^\*\|NCCL version\|^$" gpurun_out/r02_bench_aj_8gpu.err | tail -5 | cut -c1-250
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r02_bench_aj_8gpu.json"))
    print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["config"]["parallelism"])
    print({k: (v["ms_per_step"], v["value"], v["config"]["parallelism"], v["config"]["cuda_graph"]) for k, v in d["secondary"].items()})
except Exception as e:
    print("no line", e)
PY
