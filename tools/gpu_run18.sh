#!/bin/bash
# session 18: where does the cost of the folded column sums come from?  protocol only (no reads), one-warp protocol
mkdir -p gpurun_out
{
for shape in "3072 768" "768 3072"; do
  set -- $shape
  echo "--- product"; for m in 0 1; do AVN_GEMM_COLSUM=$m timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5; done
  echo "--- protocol only (8 warps wait + arrive, no reads)"; AVN_LIB_PATH=$PWD/variants/csnoread/libavion_b200.so AVN_GEMM_COLSUM=1 timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5
  echo "--- protocol only, one warp"; AVN_LIB_PATH=$PWD/variants/csonewarp/libavion_b200.so AVN_GEMM_COLSUM=1 timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5
done
} > gpurun_out/r2s18_tn_colsum_diag.txt 2>&1
cat gpurun_out/r2s18_tn_colsum_diag.txt
