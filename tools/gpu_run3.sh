#!/bin/bash
# round 2, GPU session 3: attention backward with K/V as TMEM operands - parity, A/B against the previous build, step bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -q --timeout 300 -k "attention or attn or step or block or adamw" > gpurun_out/r2s3_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s3_tests.log
tail -6 gpurun_out/r2s3_tests.log
AVN_LIB_PATH=$PWD/variants/dqred/libavion_b200.so timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 300 -k "attention or attn" > gpurun_out/r2s3_tests_dqred.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s3_tests_dqred.log
tail -3 gpurun_out/r2s3_tests_dqred.log
{
for i in 1 2; do
echo "--- new (K/V in TMEM)"; timeout 200 python tools/microbench.py 32 attn
echo "--- new + dq red"; AVN_LIB_PATH=$PWD/variants/dqred/libavion_b200.so timeout 200 python tools/microbench.py 32 attn_bwd
echo "--- base (round 1)"; AVN_LIB_PATH=$PWD/variants/r1/libavion_b200_base.so timeout 200 python tools/microbench.py 32 attn
done
echo "--- new B=256"; timeout 200 python tools/microbench.py 256 attn_bwd
echo "--- new + dq red B=256"; AVN_LIB_PATH=$PWD/variants/dqred/libavion_b200.so timeout 200 python tools/microbench.py 256 attn_bwd
echo "--- base B=256"; AVN_LIB_PATH=$PWD/variants/r1/libavion_b200_base.so timeout 200 python tools/microbench.py 256 attn_bwd
} > gpurun_out/r2s3_attn_ab.txt 2>&1
cat gpurun_out/r2s3_attn_ab.txt
timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s3_bench.json 2> gpurun_out/r2s3_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s3_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'], d['loss'])
for k in d['kernels']: print(k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
