#!/bin/bash
# round 2, GPU session 8: TN (wgrad) GEMM with the soft k-epoch barrier: parity, A/B, ncu, step bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -q --timeout 300 -k "gemm or step or block" > gpurun_out/r2s8_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s8_tests.log
tail -4 gpurun_out/r2s8_tests.log
{
for i in 1 2; do
for shape in "3072 768 803072" "768 3072 803072" "2304 768 803072" "768 768 803072"; do
  echo -n "sync on : "; timeout 120 python tools/gemm_probe.py tn $shape
  echo -n "sync off: "; AVN_TN_SYNC=0 timeout 120 python tools/gemm_probe.py tn $shape
done; done
} > gpurun_out/r2s8_tn_ab.txt 2>&1
cat gpurun_out/r2s8_tn_ab.txt
timeout 120 python tools/gemm_probe.py tn 3072 768 803072 1 > gpurun_out/r2s8_plain_tn.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -o gpurun_out/r2s8_ncu_gemm_tn_sync python tools/gemm_probe.py tn 3072 768 803072 1 > gpurun_out/r2s8_ncu_tn.log 2>&1
timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s8_bench.json 2> gpurun_out/r2s8_bench.err; echo "bench rc=$?"
AVN_TN_SYNC=0 timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s8_bench_nosync.json 2> gpurun_out/r2s8_bench_nosync.err; echo "bench rc=$?"
python - <<'PY'
import json
for n in ("r2s8_bench","r2s8_bench_nosync"):
    d=json.load(open(f'gpurun_out/{n}.json'))
    print(n, d['value'], d['ms_per_step'], d['clocks'], d['loss'])
    for k in d['kernels']: print('  ',k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
