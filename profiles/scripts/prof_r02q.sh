# round 2, GPU call Q: ncu --set full of the fused tcgen05 kernel (after tests/m2tc_time.py exited 0 without ncu in call O/P)
set -x
cd /root/repo
timeout -s KILL 120 python tests/m2tc_time.py 4096 2>&1 | tail -2
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:mlp2tc_train -s 10 -c 1 -f -o gpurun_out/r02_mlp2tc python tests/m2tc_time.py 4096 > gpurun_out/r02_ncu_q.log 2>&1; echo ncu rc=$?
tail -3 gpurun_out/r02_ncu_q.log
