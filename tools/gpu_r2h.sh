#!/bin/bash
# round-2 pass h: warp-uniform issue loops in conv_igemm / conv_wgrad + generator prefetch in the graphed step
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2h}
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -8
timeout -k 5 300 python tools/run_rooflines.py conv_igemm > gpurun_out/${T}_rooflines_conv.txt 2>&1
cut -c1-60,112-175 gpurun_out/${T}_rooflines_conv.txt
timeout -k 5 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 420 gpurun_out/${T}_bench_trainstep.json; echo; tail -3 gpurun_out/${T}_bench_trainstep.err
