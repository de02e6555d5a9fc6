# round 2, GPU call F: ncu of mlp2 v3 + sort variants at 16.7 M keys
set -x
cd /root/repo
for it in 8 16; do DCTR_SORT_ITEMS=$it python bench.py --emb-only --emb-reps 3 > gpurun_out/r02_emb_items$it.json 2>&1; done
DCTR_SORT_ITEMS=8 DCTR_SORT_BITS=8 python bench.py --emb-only --emb-reps 3 > gpurun_out/r02_emb_bits8_items8.json 2>&1
python tests/sort_bench.py > gpurun_out/r02_sort_bench_f.txt 2>&1
DCTR_MLP2=on python bench.py --steps 2 --warmup 3 --no-cpu --no-profile > gpurun_out/plain_f.log 2>&1 && \
DCTR_MLP2=on ncu --set full --clock-control none --import-source on -k regex:mlp2_train -s 3 -c 1 -o gpurun_out/r02_mlp2_v3 python bench.py --steps 2 --warmup 3 --no-cpu --no-profile > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log
