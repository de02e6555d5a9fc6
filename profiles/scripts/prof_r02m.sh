# round 2, GPU call M: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last there)
set -x
cd /root/repo
timeout -s KILL 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02_no_alloc_alone_m.txt 2>&1; echo alone rc=$?
tail -15 gpurun_out/r02_no_alloc_alone_m.txt | cut -c1-300
timeout -s KILL 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_gpu_tests_m.txt 2>&1; echo full rc=$?
tail -25 gpurun_out/r02_gpu_tests_m.txt | cut -c1-300
