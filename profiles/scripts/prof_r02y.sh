# round 2, GPU call Y (2 GPUs): row-sharded exchanges as peer-memory stores (CUDA IPC): parity scripts, then the sharded bench with phases
set -x
cd /root/repo
timeout -k 5 -s KILL 400 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_y.txt 2>&1; echo pytest rc=$?
tail -12 gpurun_out/r02_gpu_tests_y.txt | cut -c1-300
for prof in 1 ""; do
( export DCTR_SHARD_PROFILE=$prof; timeout -k 5 -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --workload fnn_criteo_sharded --steps 20 --warmup 4 --no-cpu --no-profile ) > gpurun_out/r02_bench_y_shard2.json 2> gpurun_out/r02_bench_y_shard2.err; echo bench rc=$?
grep -i "peer-memory\|Error" gpurun_out/r02_bench_y_shard2.err | head -5 | cut -c1-250
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r02_bench_y_shard2.json"))
    print(d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d.get("shard_phases_ms"))
except Exception as e:
    print("no line", e)
PY
done
