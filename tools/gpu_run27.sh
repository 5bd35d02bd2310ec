#!/bin/bash
# session 27: GEMM epilogue, residual rows fetched all at once (new) vs one unit ahead (prev), same box, interleaved
mkdir -p gpurun_out
{
for i in 1 2; do
echo "--- new"; timeout 300 python tools/microbench.py 256 gemm
echo "--- prev"; AVN_LIB_PATH=$PWD/variants/prev/libavion_b200.so timeout 300 python tools/microbench.py 256 gemm
done
} > gpurun_out/r2s27_rprefetch.txt 2>&1
grep -n "new\|prev\|res\|dgelu\|fc1 dgrad" gpurun_out/r2s27_rprefetch.txt
