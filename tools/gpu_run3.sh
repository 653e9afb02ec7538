#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -m gpu -q --tb=short --maxfail=25 -p no:cacheprovider -s > gpurun_out/r2e_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
grep -E "passed|failed|FAILED" gpurun_out/r2e_pytest.log | tail -12
VCN_BENCH_KERNELS=0 timeout -k 5 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_bench_trainstep.json 2> gpurun_out/r2e_bench_trainstep.err
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_trainstep.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')})"
VCN_CONV_PAIR=0 VCN_BENCH_KERNELS=0 timeout -k 5 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_bench_trainstep_nopair.json 2> /dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_trainstep_nopair.json')); print('nopair', {k:d[k] for k in ('value','ms_per_step','gpu_launches')})"
VCN_BENCH_KERNELS=0 timeout -k 5 300 python bench.py --workload synthesis --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_bench_synthesis.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_synthesis.json')); print('synthesis', {k:d[k] for k in ('value','ms_per_step')})"
