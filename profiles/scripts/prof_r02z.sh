# round 2, GPU call Z (8 GPUs): phases of the row-sharded step at 8 ranks (eager, CUDA events)
set -x
cd /root/repo
( export DCTR_SHARD_PROFILE=1; timeout -k 5 -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 8 --workload fnn_criteo_sharded --steps 20 --warmup 4 --no-cpu --no-profile ) > gpurun_out/r02_bench_z_shard8.json 2> gpurun_out/r02_bench_z_shard8.err; echo bench rc=$?
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r02_bench_z_shard8.json"))
    print(d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d.get("shard_phases_ms"))
except Exception as e:
    print("no line", e)
PY
