# round 2, GPU call AN: the two-deep pipelined train calls -- parity tests (tests/test_pipeline_gpu.py), then bench lines carrying
# e2e (synchronous call) and e2e_pipelined for fnn_criteo (N = 65 536) and the default workload
set -x
cd /root/repo
timeout -k 5 -s KILL 120 python -m pytest tests/test_pipeline_gpu.py -q -m gpu -x > gpurun_out/r02_gpu_tests_pipeline.txt 2>&1; echo pytest rc=$?
tail -15 gpurun_out/r02_gpu_tests_pipeline.txt | cut -c1-250
timeout -k 5 -s KILL 90 python bench.py --workload fnn_criteo --steps 50 --no-emb --no-cpu --no-secondary --no-profile > gpurun_out/r02_bench_pipe_fnn_criteo.json 2> gpurun_out/r02_bench_pipe_fnn_criteo.err; echo bench rc=$?
timeout -k 5 -s KILL 90 python bench.py --steps 200 --no-emb --no-cpu --no-secondary --no-profile > gpurun_out/r02_bench_pipe_default.json 2> gpurun_out/r02_bench_pipe_default.err; echo bench rc=$?
python - <<'PY'
import json
for f in ("fnn_criteo", "default"):
    try:
        d = json.load(open("gpurun_out/r02_bench_pipe_%s.json" % f))
        print(f, round(d["ms_per_step"], 4), d["e2e"]["ms_per_step"], d["e2e_pipelined"])
    except Exception as ex:
        print(f, "no line", ex)
PY
tail -3 gpurun_out/r02_bench_pipe_fnn_criteo.err | cut -c1-300
