#!/bin/bash
# session 19: bias gradients folded into the weight-gradient GEMM (read under the full barrier) - parity + isolated A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 600 -k "gemm" > gpurun_out/r2s19_tests_gemm.log 2>&1; echo "gemm tests rc=$?" | tee -a gpurun_out/r2s19_tests_gemm.log
tail -4 gpurun_out/r2s19_tests_gemm.log
{
for shape in "3072 768" "768 3072" "2304 768" "768 768"; do
  set -- $shape
  for m in 0 1 2; do AVN_GEMM_COLSUM=$m timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5; done
done
} > gpurun_out/r2s19_tn_colsum.txt 2>&1
cat gpurun_out/r2s19_tn_colsum.txt
