set -x
cd /root/repo
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r01.log 2>&1; echo smoke rc=$?
python bench.py > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err; echo bench rc=$?
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_r01.json 2>&1; echo ref rc=$?
python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu.log 2>&1
for k in gather_concat_kernel fold_tile_kernel sort_pass_kernel fm_fwd_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/r01_emb_$k python bench.py --emb-only --emb-reps 1 > gpurun_out/ncu_$k.log 2>&1
done
for k in gemm_tc_kernel col_wsum_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -o gpurun_out/r01_criteo_$k python bench.py --steps 1 --warmup 3 --no-cpu --no-profile --workload fnn_criteo > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out/*.ncu-rep | tail -8
