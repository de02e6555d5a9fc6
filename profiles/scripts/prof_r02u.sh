# round 2, GPU call U (2 GPUs): multi-GPU parity scripts and the default bench line at 2 GPUs with the fused tcgen05 step
set -x
cd /root/repo
timeout -s KILL 700 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_u.txt 2>&1; echo pytest rc=$?
tail -5 gpurun_out/r02_gpu_tests_u.txt | cut -c1-300
( time timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --steps 100 --warmup 5 ) > gpurun_out/r02_bench_u_dp2.json 2> gpurun_out/r02_bench_u_dp2.err; echo bench rc=$?
tail -4 gpurun_out/r02_bench_u_dp2.err | cut -c1-200
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_u_dp2.json"))
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"])
print({k: (v["ms_per_step"], v["value"]) for k, v in d["secondary"].items()})
PY
