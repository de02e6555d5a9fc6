# round 2, GPU call A: full GPU suite (new baseline-size + reference-golden tests), smoke, default bench on both GEMM paths,
# embedding sweep with Zipf and uniform ids, ncu capture of the small tensor-core GEMM of the headline step
set -x
cd /root/repo
timeout 1500 python -m pytest tests -x -q -m gpu -s > gpurun_out/r02_gpu_tests_a.txt 2>&1; echo pytest rc=$?
tail -5 gpurun_out/r02_gpu_tests_a.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_a.log 2>&1; echo smoke rc=$?
python bench.py --steps 100 > gpurun_out/r02_bench_a_tc.json 2> gpurun_out/r02_bench_a_tc.err; echo bench rc=$?
DCTR_GEMM=simt python bench.py --steps 100 --no-cpu > gpurun_out/r02_bench_a_simt.json 2> gpurun_out/r02_bench_a_simt.err; echo bench simt rc=$?
python bench.py --emb-only --emb-zipf 0 > gpurun_out/r02_emb_uniform_a.json 2>&1; echo emb rc=$?
python bench.py --steps 1 --warmup 3 --no-cpu --no-profile > gpurun_out/plain_a.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 14 -c 2 -o gpurun_out/r02_fnn_gemm_tc_small python bench.py --steps 1 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu_a.log 2>&1
ls -la gpurun_out | tail -12
