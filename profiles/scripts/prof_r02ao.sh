# round 2, GPU call AO: the whole GPU suite on the final code of the round (pipelined calls, staging refactor, bench changes)
set -x
cd /root/repo
timeout -k 5 -s KILL 200 python -m pytest tests -q -m gpu -x > gpurun_out/r02_gpu_tests_final2.txt 2>&1; echo pytest rc=$?
tail -4 gpurun_out/r02_gpu_tests_final2.txt | cut -c1-300
