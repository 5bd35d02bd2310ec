#!/bin/bash
# session 14: attention forward - share of the exponentials on the FMA pipe (AVN_FWD_POLY_NUM / DEN), interleaved A/B at B=64
mkdir -p gpurun_out
{
for i in 1 2; do
echo "--- base 1/3"; timeout 200 python tools/microbench.py 64 attn_fwd
for t in fp38 fp25 fp37 fp12; do
echo "--- $t"; AVN_LIB_PATH=$PWD/variants/$t/libavion_b200.so timeout 200 python tools/microbench.py 64 attn_fwd
done
done
} > gpurun_out/r2s14_fwd_poly.txt 2>&1
cat gpurun_out/r2s14_fwd_poly.txt
