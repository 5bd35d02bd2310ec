#!/bin/bash
# round 2, GPU session 4: attention backward B1b (P over dP, S released early) + polynomial-exp variants; ncu of LN and attn_bwd
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -q --timeout 300 -k "attention or attn or step or block" > gpurun_out/r2s4_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s4_tests.log
tail -4 gpurun_out/r2s4_tests.log
{
for i in 1 2; do
echo "--- B1b"; timeout 200 python tools/microbench.py 32 attn_bwd
echo "--- B1b poly 1/3"; AVN_LIB_PATH=$PWD/variants/poly13/libavion_b200.so timeout 200 python tools/microbench.py 32 attn_bwd
echo "--- B1b poly 1/2"; AVN_LIB_PATH=$PWD/variants/poly12/libavion_b200.so timeout 200 python tools/microbench.py 32 attn_bwd
echo "--- base (round 1)"; AVN_LIB_PATH=$PWD/variants/r1/libavion_b200_base.so timeout 200 python tools/microbench.py 32 attn_bwd
done
echo "--- B1b B=256"; timeout 200 python tools/microbench.py 256 attn_bwd
echo "--- poly13 B=256"; AVN_LIB_PATH=$PWD/variants/poly13/libavion_b200.so timeout 200 python tools/microbench.py 256 attn_bwd
echo "--- base B=256"; AVN_LIB_PATH=$PWD/variants/r1/libavion_b200_base.so timeout 200 python tools/microbench.py 256 attn_bwd
} > gpurun_out/r2s4_attn_ab.txt 2>&1
cat gpurun_out/r2s4_attn_ab.txt
timeout 200 python tools/microbench.py 64 ln > gpurun_out/r2s4_plain_ln.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ln_ -s 4 -c 2 -o gpurun_out/r2s4_ncu_ln python tools/microbench.py 64 ln > gpurun_out/r2s4_ncu_ln.log 2>&1
timeout 200 python tools/microbench.py 8 attn_bwd > gpurun_out/r2s4_plain_attn.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 2 -c 1 -o gpurun_out/r2s4_ncu_attn_bwd python tools/microbench.py 8 attn_bwd > gpurun_out/r2s4_ncu_attn.log 2>&1
ls -la gpurun_out/r2s4*.ncu-rep
