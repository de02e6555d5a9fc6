# round 2, GPU call I: compute-sanitizer memcheck over the kernels with hand-rolled look-back / tickets / mbarriers / clusters
# (sort, fold, reductions, tcgen05 GEMM, mlp2 + DSMEM reduce, dropout, arrange) on their small parity tests
set -x
cd /root/repo
SEL='test_sort_rows_bit_exact or test_scatter_sgd or test_scatter_deterministic or test_fm_scatter_and_sweep or test_arrange_fields or test_gather_concat_fwd or test_output_unit_and_reductions or (test_gemm_layouts and 100-300-177) or (test_gemm_epilogues and 130-70-45 and tanh) or test_empty_inputs'
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x -k "$SEL" > gpurun_out/r02_sanitizer_memcheck_kernels.txt 2>&1; echo memcheck kernels rc=$?
tail -8 gpurun_out/r02_sanitizer_memcheck_kernels.txt | cut -c1-200
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 python -m pytest tests/test_mlp2_gpu.py tests/test_ops_gpu.py -q -m gpu -x -k "not 4096 and not 1000" > gpurun_out/r02_sanitizer_memcheck_mlp2_ops.txt 2>&1; echo memcheck mlp2 rc=$?
tail -8 gpurun_out/r02_sanitizer_memcheck_mlp2_ops.txt | cut -c1-200
