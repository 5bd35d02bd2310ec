#!/bin/bash
# session 42: VideoMAE and ViT-L/14 single-GPU bench lines with the final kernels (bounded)
mkdir -p gpurun_out
timeout 240 python bench.py --model videomae --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s42_vmae.json 2> gpurun_out/r2s42_vmae.err; echo "vmae rc=$?"
timeout 300 python bench.py --model vitl14 --batch 112 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s42_vitl14.json 2> gpurun_out/r2s42_vitl14.err; echo "vitl14 rc=$?"
python - <<'PY'
import json
for f in ('vmae', 'vitl14'):
    try:
        d = json.load(open(f'gpurun_out/r2s42_{f}.json'))
        print(f, d['value'], d['unit'], d['ms_per_step'], d['clocks']['sm_mhz'], d['config'].get('launch'), d.get('step_tensor'))
        for k in d.get('kernels', []): print('   ', k['kernel'], round(k['ms_per_step'], 2), round(k['achieved'], 1), round(k['frac'], 3))
    except Exception as e:
        print(f, 'failed', e)
PY
tail -3 gpurun_out/r2s42_vmae.err gpurun_out/r2s42_vitl14.err
