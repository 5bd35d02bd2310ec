#!/bin/bash
# session 43 (2 GPUs): NCCL tests + CLIP bench at N=2 with the final kernels (bounded)
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_dist_gpu.py -m gpu -q -x --timeout 200 > gpurun_out/r2s43_dist_tests.log 2>&1; echo "dist tests rc=$?" | tee -a gpurun_out/r2s43_dist_tests.log
tail -3 gpurun_out/r2s43_dist_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 2 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s43_clip_n2.json 2> gpurun_out/r2s43_clip_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/r2s43_clip_n2.json'))
    print(round(d['value'], 1), round(d['ms_per_step'], 2), d['config'].get('launch'), d['clocks'], 'loss', d['loss'], 'e2e', round(d['e2e']['value'], 1))
except Exception as e:
    print('FAILED', e)
PY
tail -n 3 gpurun_out/r2s43_clip_n2.err
