#!/bin/bash
# last pass of the round: ncu --set full of the ring / direct kernels, full GPU suite, default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r3f}
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"lpips_tap_bulk|torgb_bwd|first_conv|smallc|modconv_bwd_cl_bulk" -c 14 -o /tmp/${T}_full python tools/prof_kernels.py quick lpips torgb first bwdcl > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw_ring.csv 2>/dev/null; ls -la gpurun_out/${T}_ncu_full_raw_ring.csv
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -6
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"
timeout -k 5 900 python bench.py > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 330 gpurun_out/${T}_bench_trainstep.json; echo
