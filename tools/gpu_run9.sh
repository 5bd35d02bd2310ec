#!/bin/bash
# round 2, GPU session 9: compute-sanitizer memcheck over one tiny CLIP + VideoMAE train step (every kernel family of the path)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2s9_plain.log 2>&1 && \
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2s9_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/r2s9_memcheck.log
tail -12 gpurun_out/r2s9_memcheck.log
