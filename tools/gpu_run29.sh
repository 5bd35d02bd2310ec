#!/bin/bash
# session 29: ncu of the dgrad GEMM with and without the x gelu' epilogue (fc2 dgrad shape), 128-bit R loads
mkdir -p gpurun_out
AVN_GEMM_R256=0 AVN_GEMM_EPI=dgelu timeout 120 python tools/gemm_probe.py nn 803072 3072 768 3 || exit 1
AVN_GEMM_R256=0 AVN_GEMM_EPI=dgelu timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s29_nn_dgelu python tools/gemm_probe.py nn 803072 3072 768 3 > gpurun_out/r2s29_ncu1.log 2>&1; echo "ncu dgelu rc=$?"
AVN_GEMM_R256=1 AVN_GEMM_EPI=dgelu timeout 600 ncu --set full --clock-control none -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s29_nn_dgelu256 python tools/gemm_probe.py nn 803072 3072 768 3 > gpurun_out/r2s29_ncu3.log 2>&1; echo "ncu dgelu256 rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:gemm_kernel -s 3 -c 1 -f -o gpurun_out/r2s29_nn_plain python tools/gemm_probe.py nn 803072 3072 768 3 > gpurun_out/r2s29_ncu2.log 2>&1; echo "ncu plain rc=$?"
ls -la gpurun_out/*.ncu-rep
