#!/bin/bash
# round 2, GPU session 5: packed-math LayerNorm - parity, microbench, step bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py tests/test_videomae_gpu.py -m gpu -q --timeout 300 > gpurun_out/r2s5_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s5_tests.log
tail -4 gpurun_out/r2s5_tests.log
timeout 300 python tools/microbench.py 256 ln > gpurun_out/r2s5_micro.txt 2>&1; cat gpurun_out/r2s5_micro.txt
timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s5_bench.json 2> gpurun_out/r2s5_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s5_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'], d['loss'])
for k in d['kernels']: print(k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
