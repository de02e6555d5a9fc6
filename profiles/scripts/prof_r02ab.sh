set -x
cd /root/repo
for mode in auto off; do DCTR_MLP2=$mode timeout -k 5 -s KILL 200 python -m pytest tests/test_baseline_sizes_gpu.py -q -m gpu -s -k "fnn_ipinyou" 2>&1 | grep "fnn_ipinyou step\|passed\|failed"; done
timeout -k 5 -s KILL 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_gpu_tests_final.txt 2>&1; echo full rc=$?
tail -5 gpurun_out/r02_gpu_tests_final.txt | cut -c1-300
