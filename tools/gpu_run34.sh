#!/bin/bash
# session 34: HBM write / read / copy bandwidth probe; whole GPU suite; step bench
mkdir -p gpurun_out
timeout 120 python tools/probe/hbm_write.py > gpurun_out/r2s34_hbm_write.txt 2>&1; cat gpurun_out/r2s34_hbm_write.txt
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/r2s34_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s34_tests.log
tail -3 gpurun_out/r2s34_tests.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2s34_bench.json 2> gpurun_out/r2s34_bench.err; echo "bench rc=$?"
AVN_LIB_PATH=$PWD/variants/prev/libavion_b200.so timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2s34_bench_prev.json 2> gpurun_out/r2s34_bench_prev.err; echo "bench prev rc=$?"
python - <<'PY'
import json
for f in ('', '_prev'):
    try:
        d = json.load(open(f'gpurun_out/r2s34_bench{f}.json'))
    except Exception as e:
        print(f, 'no json', e); continue
    print('bench', f, d['value'], d['ms_per_step'], d['clocks'], d['loss'], d['gpu_launches'])
    for k in d['kernels']: print('   ', k['kernel'], round(k['ms_per_step'], 2), round(k['achieved'], 1), round(k['frac'], 3))
PY
