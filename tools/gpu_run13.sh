#!/bin/bash
# session 13: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last there)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests_s13.log 2>&1; echo "suite rc=$?" >> gpurun_out/gpu_tests_s13.log
tail -3 gpurun_out/noalloc_alone.log gpurun_out/gpu_tests_s13.log
