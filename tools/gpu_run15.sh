#!/bin/bash
# session 20: bias gradients folded into the weight-gradient GEMM - parity, isolated A/B, step A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/r2s20_tests_e2e.log 2>&1; echo "e2e tests rc=$?" | tee -a gpurun_out/r2s20_tests_e2e.log
tail -4 gpurun_out/r2s20_tests_e2e.log
for f in 1 0; do
AVN_FOLD_BIAS_GRAD=$f timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2s20_bench_fold$f.json 2> gpurun_out/r2s20_bench_fold$f.err; echo "bench fold=$f rc=$?"
done
python - <<'PY'
import json
for f in (1, 0):
    try:
        d = json.load(open(f'gpurun_out/r2s20_bench_fold{f}.json'))
    except Exception as e:
        print(f, 'no json', e); continue
    print('fold', f, d['value'], d['ms_per_step'], d['clocks'], d['loss'], d['gpu_launches'])
    for k in d['kernels']: print('   ', k['kernel'], round(k['ms_per_step'], 2), round(k['achieved'], 1), round(k['frac'], 3))
PY
