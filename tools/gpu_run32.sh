#!/bin/bash
# session 32: gelu' moved from the fc1 recompute (dual) epilogue into the dgrad (dgelu) epilogue: parity + interleaved A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py tests/test_torch_ops.py tests/test_videomae_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/r2s32_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s32_tests.log
tail -3 gpurun_out/r2s32_tests.log
{
for i in 1 2 3; do
echo "--- new"; timeout 300 python tools/microbench.py 256 gemm | grep "fc1\|dgelu"
echo "--- prev"; AVN_LIB_PATH=$PWD/variants/prev/libavion_b200.so timeout 300 python tools/microbench.py 256 gemm | grep "fc1\|dgelu"
done
} > gpurun_out/r2s32_dual.txt 2>&1
cat gpurun_out/r2s32_dual.txt
