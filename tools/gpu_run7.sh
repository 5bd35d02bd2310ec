#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
AVN_LIB_PATH=$PWD/variants/trace/libavion_b200.so timeout 200 python tools/trace_attn_fwd.py 8 > gpurun_out/r2s7_trace_fwd.txt 2>&1
cat gpurun_out/r2s7_trace_fwd.txt
