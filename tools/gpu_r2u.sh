#!/bin/bash
# pass u: phase-fastest tile order of the transposed conv
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=r2u
timeout -k 5 900 python -m pytest tests/test_gpu_modconv.py tests/test_gpu_networks.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest.log
timeout -k 5 300 python tools/prof_kernels.py tconv > gpurun_out/${T}_tconv_timings.txt 2>&1; cat gpurun_out/${T}_tconv_timings.txt
timeout -k 5 300 python tools/run_rooflines.py TCONV2 > gpurun_out/${T}_rooflines_tconv.txt 2>&1; cut -c1-75,112-150 gpurun_out/${T}_rooflines_tconv.txt
export VCN_BENCH_KERNELS=0
for i in 1 2; do timeout -k 5 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null | head -c 260; echo; done
