#!/bin/bash
# session 50: attention forward - head start of softmax group 0 over group 1 (AVN_FWD_LAG cycles; shipped: 1000); bounded
mkdir -p gpurun_out
{
for i in 1 2; do
echo "--- lag 1000 (shipped)"; timeout 60 python tools/microbench.py 64 attn_fwd
for l in 700 1200 1400 1700; do
echo "--- lag $l"; AVN_LIB_PATH=$PWD/variants/lag$l/libavion_b200.so timeout 60 python tools/microbench.py 64 attn_fwd
done
done
} > gpurun_out/r2s50_fwd_lag.txt 2>&1
cat gpurun_out/r2s50_fwd_lag.txt
