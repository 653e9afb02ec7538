#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export VCN_BENCH_GRAPH=0 VCN_BENCH_KERNELS=0
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/r2c_launches_trainstep.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2c_ncu.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/r2c_ncu.log; wc -l gpurun_out/r2c_launches_trainstep.csv
