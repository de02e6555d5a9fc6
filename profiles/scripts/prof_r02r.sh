# round 2, GPU call R: whole GPU suite with the fused tcgen05 step as the default for two-hidden-layer nets
set -x
cd /root/repo
timeout -s KILL 1500 python -m pytest tests -q -m gpu > gpurun_out/r02_gpu_tests_r.txt 2>&1; echo full rc=$?
tail -40 gpurun_out/r02_gpu_tests_r.txt | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
