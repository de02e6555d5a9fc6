cd /root/repo
python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r01b.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 6 -c 1 -o gpurun_out/r01b_criteo_gemm_tc_kernel python bench.py --steps 1 --warmup 3 --no-cpu --no-profile --workload fnn_criteo > gpurun_out/ncu_gemm2.log 2>&1
ls -la gpurun_out/r01b*
