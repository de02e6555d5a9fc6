# round 2, GPU call N (2 GPUs): multi-GPU parity scripts inside pytest with the fixed launcher
set -x
cd /root/repo
nvidia-smi -L
timeout -s KILL 700 python -m pytest tests/test_multi_gpu.py -q -m gpu -s > gpurun_out/r02_gpu_tests_n.txt 2>&1; echo pytest rc=$?
tail -25 gpurun_out/r02_gpu_tests_n.txt | cut -c1-300
