#!/bin/bash
# session 48: GEMM A-operand loads with an L2 evict_last hint (variant evlast) vs plain (dual stores evict_first in both); bounded
mkdir -p gpurun_out
AVN_LIB_PATH=$PWD/variants/evlast/libavion_b200.so timeout 150 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 100 -k "gemm" > gpurun_out/r2s48_tests.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/r2s48_tests.log
{
for i in 1 2; do
echo "--- plain"; timeout 90 python tools/microbench.py 256 gemm | grep "nt\|dgelu\|dgrad"
echo "--- A loads evict_last"; AVN_LIB_PATH=$PWD/variants/evlast/libavion_b200.so timeout 90 python tools/microbench.py 256 gemm | grep "nt\|dgelu\|dgrad"
done
} > gpurun_out/r2s48_evict.txt 2>&1
cat gpurun_out/r2s48_evict.txt
