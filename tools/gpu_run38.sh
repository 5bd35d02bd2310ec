#!/bin/bash
# session 38: persistent attention forward (original exponential / P-store order, full register budget): parity + A/B, tightly bounded
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x --timeout 100 -k "attn or attention" > gpurun_out/r2s38_tests.log 2>&1; rc=$?; echo "attn tests rc=$rc" | tee -a gpurun_out/r2s38_tests.log
tail -3 gpurun_out/r2s38_tests.log
if [ $rc -ne 0 ]; then exit 1; fi
{
for i in 1 2; do
echo "--- persistent"; timeout 60 python tools/microbench.py 64 attn_fwd; timeout 60 python tools/microbench.py 256 attn_fwd
echo "--- one CTA per item"; AVN_ATTN_FWD_PERSIST=0 timeout 60 python tools/microbench.py 64 attn_fwd; AVN_ATTN_FWD_PERSIST=0 timeout 60 python tools/microbench.py 256 attn_fwd
done
} > gpurun_out/r2s38_fwd_persist.txt 2>&1
cat gpurun_out/r2s38_fwd_persist.txt
