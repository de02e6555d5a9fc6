# round 2, GPU call B: whole GPU suite (fused small-net kernels, double-precision decay in the fold), smoke, default bench
set -x
cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/r02_gpu_tests_b.txt 2>&1; echo pytest rc=$?
tail -15 gpurun_out/r02_gpu_tests_b.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_b.log 2>&1; echo smoke rc=$?; tail -3 gpurun_out/r02_smoke_b.log
python bench.py --steps 100 --no-cpu > gpurun_out/r02_bench_b.json 2> gpurun_out/r02_bench_b.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_b.err
