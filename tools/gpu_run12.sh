#!/bin/bash
# round 2, GPU session 12: transform kernel with per-thread pixel pairs; full GPU suite; reference arm; default bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r2s12_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s12_tests.log
tail -4 gpurun_out/r2s12_tests.log
timeout 300 python tools/microbench.py 256 transform > gpurun_out/r2s12_micro.txt 2>&1; cat gpurun_out/r2s12_micro.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2s12_ref.json 2> gpurun_out/r2s12_ref.err; echo "ref rc=$?"; head -c 900 gpurun_out/r2s12_ref.json; echo
timeout 600 python bench.py --steps 8 --warmup 3 > gpurun_out/r2s12_bench.json 2> gpurun_out/r2s12_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s12_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'], d['loss'], 'e2e', d['e2e']['value'], d['cpu_baseline'])
for k in d['kernels']: print('  ',k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
