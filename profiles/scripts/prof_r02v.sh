# round 2, GPU call V (2 GPUs): fixed-capacity row-sharded step: unit test, parity scripts, bench of the sharded workload
set -x
cd /root/repo
timeout -s KILL 300 python -m pytest tests/test_ops_gpu.py -q -m gpu -k shard_pack 2>&1 | tail -3 | cut -c1-300
timeout -s KILL 700 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_v.txt 2>&1; echo pytest rc=$?
tail -8 gpurun_out/r02_gpu_tests_v.txt | cut -c1-300
for prof in 1 ""; do
( export DCTR_SHARD_PROFILE=$prof; timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --workload fnn_criteo_sharded --steps 20 --warmup 4 --no-cpu --no-profile ) > gpurun_out/r02_bench_v_shard2.json 2> gpurun_out/r02_bench_v_shard2.err; echo bench rc=$?
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_v_shard2.json"))
print(d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["gpu_launches"], d.get("shard_phases_ms"))
PY
done
