# round 2, GPU call H (2 GPUs): multi-GPU parity scripts inside pytest, the default bench line at 2 GPUs with its secondary block
set -x
cd /root/repo
nvidia-smi -L
timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_h.txt 2>&1; echo pytest rc=$?
tail -25 gpurun_out/r02_gpu_tests_h.txt | cut -c1-300
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29570 bench.py --gpus 2 --steps 100 --warmup 5 ) > gpurun_out/r02_bench_h_dp2.json 2> gpurun_out/r02_bench_h_dp2.err; echo bench rc=$?
tail -6 gpurun_out/r02_bench_h_dp2.err | cut -c1-300
