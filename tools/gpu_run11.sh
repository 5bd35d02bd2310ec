#!/bin/bash
# round 2, GPU session 11: ncu launch list of the default bench command (kernel shares), full captures of the memory-bound kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2s11_plain_bench.json 2> gpurun_out/r2s11_plain_bench.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2s11_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r2s11_ncu_bench.log 2>&1
echo "launch list rc=$?"
timeout 200 python tools/microbench.py 64 ln > gpurun_out/r2s11_plain_ln.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ln_ -s 4 -c 2 -o gpurun_out/r2s11_ncu_ln python tools/microbench.py 64 ln > gpurun_out/r2s11_ncu_ln.log 2>&1
timeout 200 python tools/microbench.py 64 transform > gpurun_out/r2s11_plain_tr.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:clip_ -s 2 -c 4 -o gpurun_out/r2s11_ncu_transform python tools/microbench.py 64 transform > gpurun_out/r2s11_ncu_tr.log 2>&1
ls -la gpurun_out/r2s11*
gzip -f gpurun_out/r2s11_launches.csv
