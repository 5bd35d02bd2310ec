#!/bin/bash
# session 23: MN-major operands staged by one 4-D TMA box per stage (AVN_GEMM_BOX4=1) vs one 3-D box per 64-column chunk (=0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 600 -k "gemm" > gpurun_out/r2s23_tests_gemm.log 2>&1; echo "gemm tests rc=$?" | tee -a gpurun_out/r2s23_tests_gemm.log
tail -3 gpurun_out/r2s23_tests_gemm.log
{
for i in 1 2; do
for b in 0 1; do
  for shape in "3072 768" "768 3072" "2304 768" "768 768"; do set -- $shape
    AVN_GEMM_BOX4=$b AVN_GEMM_COLSUM=1 timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5 | sed "s/^/box4=$b /"
  done
  for shape in "768 2304" "768 3072" "3072 768" "768 768"; do set -- $shape
    AVN_GEMM_BOX4=$b timeout 120 python tools/gemm_probe.py nn 803072 $1 $2 5 | sed "s/^/box4=$b /"
  done
done
done
} > gpurun_out/r2s23_box4.txt 2>&1
cat gpurun_out/r2s23_box4.txt
