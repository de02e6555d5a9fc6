# round 2, GPU call E: torch.ops boundary in the engines, mlp2 v3 (loop bodies that fit the instruction cache, 64x64 wgrad tiles)
set -x
cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/r02_gpu_tests_e.txt 2>&1; echo pytest rc=$?
grep -E "passed|failed|FAILED|Error" gpurun_out/r02_gpu_tests_e.txt | tail -15
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_e.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r02_smoke_e.log
python bench.py --steps 100 --no-cpu > gpurun_out/r02_bench_e.json 2> gpurun_out/r02_bench_e.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_e.err
DCTR_MLP2=off python bench.py --steps 100 --no-cpu --no-profile > gpurun_out/r02_bench_e_off.json 2>&1; echo rc=$?
python bench.py --steps 50 --no-cpu --no-profile --workload fnn_criteo > gpurun_out/r02_bench_e_criteo.json 2>&1; echo rc=$?
