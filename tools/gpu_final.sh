#!/bin/bash
# final check of the committed state: full GPU suite + the default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2zz}
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -6
timeout -k 5 300 python tools/run_rooflines.py first_conv > gpurun_out/${T}_rooflines_first.txt 2>&1; cut -c1-75,112-150 gpurun_out/${T}_rooflines_first.txt
timeout -k 5 900 python bench.py > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 330 gpurun_out/${T}_bench_trainstep.json; echo
