#!/bin/bash
# session 30: residual / gelu' tile staged by TMA (AVN_GEMM_RTMA=1) vs row-per-thread global loads (=0): parity, interleaved A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "gemm" 2>&1 | tail -3
AVN_GEMM_RTMA=0 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "gemm" 2>&1 | tail -2
{
for i in 1 2 3; do
for v in 1 0; do
echo "--- rtma=$v"; AVN_GEMM_RTMA=$v timeout 300 python tools/microbench.py 256 gemm | grep "res\|dgelu\|fc1 dgrad\|qkv bias"
done
done
} > gpurun_out/r2s30_rtma.txt 2>&1
cat gpurun_out/r2s30_rtma.txt
