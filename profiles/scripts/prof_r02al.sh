# round 2, GPU call AL (last 4.9 GPU-minutes of the round): ncu --set full of the CURRENT fold tile kernel at the embedding-sweep size
# (the traffic literal in bench.py was a round-1 capture of the kernel before the update list / double-precision decay).
# bench.py --emb-only exited 0 without ncu on this code in call AK.  Launches of fold_tile_kernel with --emb-reps 1: two per variant
# (warm-up + timed) in the order seq_decay, sum, masked seq_decay, ...: -s 2 -c 3 = sum, sum, masked seq_decay.
set -x
cd /root/repo
timeout -k 5 -s KILL 170 ncu --set full --clock-control none --import-source off -k regex:fold_tile_kernel -s 2 -c 3 -f -o gpurun_out/r02_fold_final python bench.py --emb-only --emb-reps 1 > gpurun_out/r02_ncu_al.log 2>&1; echo ncu rc=$?
tail -3 gpurun_out/r02_ncu_al.log | cut -c1-300
ls -la gpurun_out/*.ncu-rep
