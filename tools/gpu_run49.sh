#!/bin/bash
# session 49: committed tree (dual epilogue with evict-first stores) - whole GPU suite + default bench (bounded)
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x --timeout 150 > gpurun_out/r2s49_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s49_tests.log
tail -3 gpurun_out/r2s49_tests.log
timeout 300 python bench.py --steps 8 --warmup 3 > gpurun_out/r2s49_bench.json 2> gpurun_out/r2s49_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s49_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'], d['loss'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
for k in d['kernels']: print('  ',k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
