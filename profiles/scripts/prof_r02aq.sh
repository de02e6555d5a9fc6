# round 2, GPU call AQ (last GPU-seconds): the reference arm on the GPU box's host cores, with the reference's own script (B0) beside the port
set -x
cd /root/repo
( time timeout -k 5 -s KILL 40 python bench.py --impl reference --steps 10 --warmup 3 ) > gpurun_out/r02_bench_reference_b0.json 2> gpurun_out/r02_bench_reference_b0.err; echo rc=$?
cut -c1-1500 gpurun_out/r02_bench_reference_b0.json
tail -5 gpurun_out/r02_bench_reference_b0.err | cut -c1-200
nproc
