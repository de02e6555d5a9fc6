# round 2, GPU call AG (2 GPUs): bucket-order gradient push + dense all-reduce next to the fold: op test, 2-rank parity scripts, 2-GPU bench line
set -x
cd /root/repo
timeout -k 5 -s KILL 200 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "slot_grads or shard_pack" > gpurun_out/r02_gpu_tests_ag_ops.txt 2>&1; echo ops rc=$?
tail -3 gpurun_out/r02_gpu_tests_ag_ops.txt | cut -c1-300
timeout -k 5 -s KILL 500 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_ag.txt 2>&1; echo pytest rc=$?
tail -5 gpurun_out/r02_gpu_tests_ag.txt | cut -c1-300
timeout -k 5 -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/r02_bench_ag_2gpu.json 2> gpurun_out/r02_bench_ag_2gpu.err; echo bench rc=$?
tail -4 gpurun_out/r02_bench_ag_2gpu.err | cut -c1-200
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_ag_2gpu.json"))
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"])
print({k: (v["ms_per_step"], v["value"]) for k, v in d["secondary"].items()})
PY
