#!/bin/bash
# round 2, GPU session 1: parity tests, same-box comparators, default bench, smoke
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2s1_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r2s1_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2s1_tests.log
tail -5 gpurun_out/r2s1_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2s1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2s1_smoke.log
tail -3 gpurun_out/r2s1_smoke.log
timeout 600 python tools/compare_same_box.py --out gpurun_out/r2s1_same_box.json > gpurun_out/r2s1_same_box.txt 2>&1; echo "cmp rc=$?" >> gpurun_out/r2s1_same_box.txt
tail -40 gpurun_out/r2s1_same_box.txt
timeout 400 python bench.py --steps 6 --warmup 3 > gpurun_out/r2s1_bench.json 2> gpurun_out/r2s1_bench.err; echo "bench rc=$?"
cat gpurun_out/r2s1_bench.json | head -c 3000
