# round 2, GPU call G: GPU suite (in-graph dropout, bench refactor), the full default bench line incl. the secondary block and both
# embedding sweeps, the reference arm
set -x
cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/r02_gpu_tests_g.txt 2>&1; echo pytest rc=$?
grep -E "passed|failed|FAILED|Error" gpurun_out/r02_gpu_tests_g.txt | tail -12
( time python bench.py ) > gpurun_out/r02_bench_g.json 2> gpurun_out/r02_bench_g.err; echo bench rc=$?; tail -5 gpurun_out/r02_bench_g.err
( time python bench.py --impl reference --steps 20 --warmup 2 ) > gpurun_out/r02_bench_g_ref.json 2> gpurun_out/r02_bench_g_ref.err; echo rc=$?; tail -4 gpurun_out/r02_bench_g_ref.err
