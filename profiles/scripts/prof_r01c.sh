cd /root/repo
python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r01c.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sort_pass_kernel -s 4 -c 2 -o gpurun_out/r01c_fnn_sort_pass python bench.py --steps 1 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu_sp.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sort_hist_kernel -s 2 -c 1 -o gpurun_out/r01c_fnn_sort_hist python bench.py --steps 1 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu_sh.log 2>&1
ls -la gpurun_out/r01c*
