#!/bin/bash
# session 39: final state - whole GPU suite, default bench (with cpu baseline), smoke, ncu launch list of the step
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x --timeout 150 > gpurun_out/r2s39_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s39_tests.log
tail -3 gpurun_out/r2s39_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2s39_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2s39_smoke.txt
timeout 300 python bench.py --steps 8 --warmup 3 > gpurun_out/r2s39_bench.json 2> gpurun_out/r2s39_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s39_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'], d['loss'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], d['roofline'])
for k in d['kernels']: print('  ',k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
timeout 420 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2s39_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2s39_ncu_bench.log 2>&1; echo "ncu rc=$?"
gzip -f gpurun_out/r2s39_launches.csv
ls -la gpurun_out/r2s39_*
