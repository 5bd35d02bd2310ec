#!/bin/bash
# session 46: last sanity of the committed tree - whole GPU suite + smoke (bounded)
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x --timeout 150 > gpurun_out/r2s46_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2s46_tests.log
tail -3 gpurun_out/r2s46_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | head -c 600; echo
