#!/bin/bash
# session 51: ncu --set full of the final attention backward (persistent) and of the wgrad GEMM (4-D boxes + bias gradient); bounded
mkdir -p gpurun_out
timeout 60 python tools/microbench.py 32 attn_bwd || exit 1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 2 -c 1 -f -o gpurun_out/r2s51_attn_bwd python tools/microbench.py 32 attn_bwd > gpurun_out/r2s51_ncu1.log 2>&1; echo "ncu attn_bwd rc=$?"
AVN_GEMM_COLSUM=1 timeout 200 ncu --set full --clock-control none -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s51_gemm_tn python tools/gemm_probe.py tn 3072 768 803072 3 > gpurun_out/r2s51_ncu2.log 2>&1; echo "ncu gemm_tn rc=$?"
ls -la gpurun_out/r2s51_*.ncu-rep
