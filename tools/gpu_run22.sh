#!/bin/bash
# session 22: wgrad GEMM - L2 prefetch distance of the producer (AVN_GEMM_TN_L2_AHEAD k-blocks)
mkdir -p gpurun_out
{
for d in 0 4 8 16 32 64; do
  for shape in "3072 768" "768 3072" "2304 768" "768 768"; do set -- $shape
    AVN_GEMM_TN_L2_AHEAD=$d AVN_GEMM_COLSUM=1 timeout 120 python tools/gemm_probe.py tn $1 $2 803072 5 | sed "s/^/l2_ahead=$d /"
  done
done
} > gpurun_out/r2s22_tn_l2ahead.txt 2>&1
cat gpurun_out/r2s22_tn_l2ahead.txt
