#!/bin/bash
# session 33: ncu of the fc1 recompute GEMM (dual epilogue) and of bias+gelu
mkdir -p gpurun_out
AVN_GEMM_EPI=dual timeout 120 python tools/gemm_probe.py nt 803072 3072 768 3 || exit 1
AVN_GEMM_EPI=dual timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s33_nt_dual python tools/gemm_probe.py nt 803072 3072 768 3 > gpurun_out/r2s33_ncu1.log 2>&1; echo "ncu dual rc=$?"
AVN_GEMM_EPI=gelu timeout 600 ncu --set full --clock-control none -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s33_nt_gelu python tools/gemm_probe.py nt 803072 3072 768 3 > gpurun_out/r2s33_ncu2.log 2>&1; echo "ncu gelu rc=$?"
