#!/bin/bash
# session 37: persistent attention forward (+ backward with its full register budget): parity, A/B persistent vs one CTA per item
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 300 -k "attn or attention" > gpurun_out/r2s37_tests.log 2>&1; echo "attn tests rc=$?" | tee -a gpurun_out/r2s37_tests.log
tail -3 gpurun_out/r2s37_tests.log
AVN_ATTN_FWD_PERSIST=0 AVN_ATTN_BWD_PERSIST=0 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 300 -k "attn or attention" 2>&1 | tail -2
timeout 900 python -m pytest tests/test_e2e_gpu.py tests/test_videomae_gpu.py tests/test_torch_ops.py -m gpu -q -x --timeout 600 2>&1 | tail -2
{
for i in 1 2; do
echo "--- persistent"; timeout 200 python tools/microbench.py 64 attn; timeout 200 python tools/microbench.py 256 attn
echo "--- one CTA per item"; AVN_ATTN_FWD_PERSIST=0 AVN_ATTN_BWD_PERSIST=0 timeout 200 python tools/microbench.py 64 attn; AVN_ATTN_FWD_PERSIST=0 AVN_ATTN_BWD_PERSIST=0 timeout 200 python tools/microbench.py 256 attn
done
} > gpurun_out/r2s37_attn_persist.txt 2>&1
cat gpurun_out/r2s37_attn_persist.txt
