#!/bin/bash
# session 28: residual / gelu' rows read with 256-bit loads (AVN_GEMM_R256=1) vs 128-bit (=0); interleaved, three rounds
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k gemm 2>&1 | tail -2
{
for i in 1 2 3; do
for v in 1 0; do
echo "--- r256=$v"; AVN_GEMM_R256=$v timeout 300 python tools/microbench.py 256 gemm | grep "res\|dgelu\|fc1 dgrad"
done
done
} > gpurun_out/r2s28_r256.txt 2>&1
cat gpurun_out/r2s28_r256.txt
