#!/bin/bash
# round-2 pass k: small-C ToRGB backward pack, bulk-staged LPIPS taps, VGG conv1_1 on the implicit GEMM, 3x3/s2 max-pool
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r2k
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -8
timeout -k 5 400 python tools/run_rooflines.py > gpurun_out/${T}_rooflines.txt 2>&1
cut -c1-75,112-150 gpurun_out/${T}_rooflines.txt
timeout -k 5 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 300 gpurun_out/${T}_bench_trainstep.json; echo; tail -3 gpurun_out/${T}_bench_trainstep.err
