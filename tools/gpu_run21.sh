#!/bin/bash
# session 21: VideoMAE / classifier tests after folding their bias gradients; CTA-exit TMA wait (.read vs full) A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_videomae_gpu.py tests/test_classifier.py tests/test_e2e_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/r2s21_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s21_tests.log
tail -3 gpurun_out/r2s21_tests.log
AVN_LIB_PATH=$PWD/variants/exitread/libavion_b200.so timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 600 -k "attn or attention or gemm" > gpurun_out/r2s21_tests_exitread.log 2>&1; echo "exitread tests rc=$?" | tee -a gpurun_out/r2s21_tests_exitread.log
tail -3 gpurun_out/r2s21_tests_exitread.log
{
for i in 1 2; do
echo "--- base"; timeout 200 python tools/microbench.py 64 attn
echo "--- exit .read"; AVN_LIB_PATH=$PWD/variants/exitread/libavion_b200.so timeout 200 python tools/microbench.py 64 attn
done
echo "--- base"; timeout 100 python tools/gemm_probe.py nt 803072 3072 768; timeout 100 python tools/gemm_probe.py tn 3072 768 803072
echo "--- exit .read"; AVN_LIB_PATH=$PWD/variants/exitread/libavion_b200.so timeout 100 python tools/gemm_probe.py nt 803072 3072 768; AVN_LIB_PATH=$PWD/variants/exitread/libavion_b200.so timeout 100 python tools/gemm_probe.py tn 3072 768 803072
} > gpurun_out/r2s21_exitread.txt 2>&1
cat gpurun_out/r2s21_exitread.txt
