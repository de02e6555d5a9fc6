# round 2, GPU call AM: K6 (sort + fold) as one global list vs L2-blocked (tests/k6_blocked_bench.py)
set -x
cd /root/repo
timeout -k 5 -s KILL 150 python tests/k6_blocked_bench.py 5 > gpurun_out/r02_k6_blocked.json 2> gpurun_out/r02_k6_blocked.err; echo rc=$?
grep -v "^+" gpurun_out/r02_k6_blocked.err | tail -12
