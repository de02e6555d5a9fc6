# round 2, GPU call D: ncu --set full of the fused small-net kernels (why 200 us?)
set -x
cd /root/repo
python bench.py --steps 2 --warmup 3 --no-cpu --no-profile > gpurun_out/plain_d.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:mlp2_ -s 9 -c 3 -o gpurun_out/r02_mlp2_v2 python bench.py --steps 2 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu_d.log 2>&1
tail -3 gpurun_out/ncu_d.log
