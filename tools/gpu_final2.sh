#!/bin/bash
# final check: full GPU suite, default bench line, the two other workloads
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r3c}
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -6
timeout -k 5 900 python bench.py > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 330 gpurun_out/${T}_bench_trainstep.json; echo
timeout -k 5 300 python bench.py --workload opsweep --steps 20 --warmup 3 > gpurun_out/${T}_bench_opsweep.json 2> gpurun_out/${T}_bench_opsweep.err; echo "opsweep rc=$?"; head -c 260 gpurun_out/${T}_bench_opsweep.json; echo; tail -2 gpurun_out/${T}_bench_opsweep.err
VCN_BENCH_GRAPH=0 VCN_BENCH_KERNELS=0 timeout -k 5 300 python bench.py --workload opsweep --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | head -c 260; echo
timeout -k 5 300 python bench.py --workload synthesis --steps 20 --warmup 3 > gpurun_out/${T}_bench_synthesis.json 2>/dev/null; head -c 200 gpurun_out/${T}_bench_synthesis.json; echo
