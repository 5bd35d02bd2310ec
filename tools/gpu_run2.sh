#!/bin/bash
# round 2, GPU session 2: full GPU test suite with the TMA LayerNorm + new transform, LN / transform microbench, TN GEMM sweep + ncu
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r2s2_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s2_tests.log
tail -15 gpurun_out/r2s2_tests.log
{ timeout 300 python tools/microbench.py 256 ln; timeout 300 python tools/microbench.py 256 transform; } > gpurun_out/r2s2_micro.txt 2>&1
cat gpurun_out/r2s2_micro.txt
{
for shape in "3072 768 803072" "768 3072 803072" "2304 768 803072" "768 768 803072"; do
  timeout 120 python tools/gemm_probe.py tn $shape
  AVN_GEMM_CG=1 timeout 120 python tools/gemm_probe.py tn $shape
  for sp in 1 2 4 6; do AVN_GEMM_SPLITS=$sp timeout 120 python tools/gemm_probe.py tn $shape; done
done
timeout 120 python tools/gemm_probe.py nt 803072 3072 768
timeout 120 python tools/gemm_probe.py nn 803072 768 3072
} > gpurun_out/r2s2_gemm_sweep.txt 2>&1
cat gpurun_out/r2s2_gemm_sweep.txt
timeout 120 python tools/gemm_probe.py tn 3072 768 803072 1 > gpurun_out/r2s2_plain_tn.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -o gpurun_out/r2s2_ncu_gemm_tn python tools/gemm_probe.py tn 3072 768 803072 1 > gpurun_out/r2s2_ncu_tn.log 2>&1
timeout 120 python tools/gemm_probe.py nt 803072 3072 768 1 > gpurun_out/r2s2_plain_nt.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -o gpurun_out/r2s2_ncu_gemm_nt python tools/gemm_probe.py nt 803072 3072 768 1 > gpurun_out/r2s2_ncu_nt.log 2>&1
ls -la gpurun_out/*.ncu-rep
timeout 400 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2s2_bench.json 2> gpurun_out/r2s2_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s2_bench.json'))
print(d['value'], d['ms_per_step'], d['clocks'])
for k in d['kernels']: print(k['kernel'], round(k['ms_per_step'],2), round(k['achieved'],1), round(k['frac'],3))
PY
