#!/bin/bash
# session 31: whole GPU suite + step bench with the TMA-staged R tile (AVN_GEMM_RTMA 1 / 0 back to back)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/r2s31_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s31_tests.log
tail -3 gpurun_out/r2s31_tests.log
for b in 1 0; do
AVN_GEMM_RTMA=$b timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2s31_bench_rtma$b.json 2> gpurun_out/r2s31_bench_rtma$b.err; echo "bench rtma=$b rc=$?"
done
python - <<'PY'
import json
for f in (1, 0):
    try:
        d = json.load(open(f'gpurun_out/r2s31_bench_rtma{f}.json'))
    except Exception as e:
        print(f, 'no json', e); continue
    print('rtma', f, d['value'], d['ms_per_step'], d['clocks'], d['loss'], d['gpu_launches'])
    for k in d['kernels']: print('   ', k['kernel'], round(k['ms_per_step'], 2), round(k['achieved'], 1), round(k['frac'], 3))
PY
