# round 2, GPU call T: ncu --set full of the final fused kernel (tests/m2tc_time.py ran plain with exit 0 first)
set -x
cd /root/repo
timeout -s KILL 120 python tests/m2tc_time.py 4096 2>&1 | tail -2 | tee gpurun_out/r02_m2tc_time_t.txt; echo plain rc=$?
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:mlp2tc_train -s 10 -c 1 -f -o gpurun_out/r02_mlp2tc_final python tests/m2tc_time.py 4096 > gpurun_out/r02_ncu_t.log 2>&1; echo ncu rc=$?
