#!/bin/bash
# session 44 (8 GPUs): BASELINE.json configs[2] with the final kernels (one bounded run)
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 8 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s44_clip_n8.json 2> gpurun_out/r2s44_clip_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/r2s44_clip_n8.json'))
    print(round(d['value'], 1), round(d['ms_per_step'], 2), d['config'].get('launch'), d['clocks'], 'loss', d['loss'], 'e2e', round(d['e2e']['value'], 1))
except Exception as e:
    print('FAILED', e)
PY
tail -n 2 gpurun_out/r2s44_clip_n8.err
