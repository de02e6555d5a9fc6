# round 2, GPU call C: GPU suite with the 2-D register-tile mlp2 kernels + cluster/DSMEM weight gradients, bench, sort variants
set -x
cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu -s > gpurun_out/r02_gpu_tests_c.txt 2>&1; echo pytest rc=$?
grep -E "passed|failed|FAILED" gpurun_out/r02_gpu_tests_c.txt | tail -12
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_c.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r02_smoke_c.log
python bench.py --steps 100 --no-cpu > gpurun_out/r02_bench_c.json 2> gpurun_out/r02_bench_c.err; echo bench rc=$?; tail -3 gpurun_out/r02_bench_c.err
python bench.py --steps 100 --no-cpu --no-profile --workload fm_ipinyou > gpurun_out/r02_bench_c_fm.json 2>&1; echo rc=$?
for v in items4 items16; do DCTR_LIB=/root/repo/deep-ctr_b200/libdctr_$v.so python bench.py --emb-only --emb-reps 3 > gpurun_out/r02_emb_$v.json 2>&1; done
DCTR_SORT_BITS=9 python bench.py --emb-only --emb-reps 3 > gpurun_out/r02_emb_bits9.json 2>&1
DCTR_SORT_BITS=9 DCTR_LIB=/root/repo/deep-ctr_b200/libdctr_items16.so python bench.py --emb-only --emb-reps 3 > gpurun_out/r02_emb_bits9_items16.json 2>&1
python tests/sort_bench.py > gpurun_out/r02_sort_bench_c.txt 2>&1
