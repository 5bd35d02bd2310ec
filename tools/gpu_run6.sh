#!/bin/bash
# round 2, GPU session 6 (2 GPUs): NCCL tests of every ClipLoss mode, graphed CLIP / VideoMAE steps over 2 ranks vs eager DDP
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout 400 > gpurun_out/r2s6_dist_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s6_dist_tests.log
tail -4 gpurun_out/r2s6_dist_tests.log
run2() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; echo "$2 rc=$?"; }
run2 29511 r2s6_vmae_n2_graph --model videomae --steps 20 --warmup 5 --no-cpu-baseline
run2 29512 r2s6_vmae_n2_eager --model videomae --steps 20 --warmup 5 --no-cpu-baseline --no-graph
run2 29513 r2s6_clip_n2_eager --steps 4 --warmup 3 --no-cpu-baseline
run2 29514 r2s6_clip_n2_graph --steps 4 --warmup 3 --no-cpu-baseline --graph on
python - <<'PY'
import json
for n in ("r2s6_vmae_n2_graph","r2s6_vmae_n2_eager","r2s6_clip_n2_eager","r2s6_clip_n2_graph"):
    try:
        d=json.load(open(f'gpurun_out/{n}.json'))
        print(n, round(d['value'],1), round(d['ms_per_step'],2), d['config'].get('launch'), d['clocks']['sm_mhz'], 'loss', d['loss'], 'e2e', round(d['e2e']['value'],1))
    except Exception as e:
        print(n, 'FAILED', e)
PY
tail -3 gpurun_out/r2s6_clip_n2_graph.err
