#!/bin/bash
# round 2, GPU session 10 (8 GPUs): BASELINE.json configs[2], [3], [4] as stated
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run8() { timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; echo "$2 rc=$?"; }
run8 29611 r2s10_vmae_n8_graph --model videomae --steps 30 --warmup 5 --no-cpu-baseline
run8 29612 r2s10_vitl14_b112_n8 --model vitl14 --batch 112 --steps 4 --warmup 3 --no-cpu-baseline
run8 29613 r2s10_clip_n8 --steps 6 --warmup 3 --no-cpu-baseline
python - <<'PY'
import json
for n in ("r2s10_vmae_n8_graph","r2s10_vitl14_b112_n8","r2s10_clip_n8"):
    try:
        d=json.load(open(f'gpurun_out/{n}.json'))
        print(n, round(d['value'],1), round(d['ms_per_step'],2), d['config'].get('launch'), d['clocks'], 'loss', d['loss'], 'e2e', round(d['e2e']['value'],1), 'step_tf', d['step_tensor'])
    except Exception as e:
        print(n, 'FAILED', e)
PY
tail -3 gpurun_out/r2s10_vitl14_b112_n8.err
