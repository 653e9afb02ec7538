#!/bin/bash
# round-2 final evidence pass: full GPU suite, bench lines of every workload / arm, launch list, ncu --set full raw pages
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=r2z
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -6
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${T}_smoke.log
timeout -k 5 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 330 gpurun_out/${T}_bench_trainstep.json; echo
timeout -k 5 300 python bench.py --workload synthesis --steps 20 --warmup 3 > gpurun_out/${T}_bench_synthesis.json 2>/dev/null; head -c 200 gpurun_out/${T}_bench_synthesis.json; echo
timeout -k 5 300 python bench.py --workload opsweep --steps 20 --warmup 3 > gpurun_out/${T}_bench_opsweep.json 2>/dev/null; head -c 200 gpurun_out/${T}_bench_opsweep.json; echo
timeout -k 5 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2>/dev/null; head -c 260 gpurun_out/${T}_bench_reference_arm.json; echo
timeout -k 5 600 python bench.py --impl reference-gpu --steps 5 --warmup 3 > gpurun_out/${T}_bench_trainstep_refgpu.json 2> gpurun_out/${T}_bench_trainstep_refgpu.err; head -c 260 gpurun_out/${T}_bench_trainstep_refgpu.json; echo; tail -2 gpurun_out/${T}_bench_trainstep_refgpu.err
export VCN_BENCH_GRAPH=0 VCN_BENCH_KERNELS=0
timeout -k 5 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain.log 2>&1 &&
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/${T}_launches_trainstep.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
echo "launchlist rc=$?"; wc -l gpurun_out/${T}_launches_trainstep.csv
python tools/launch_summary.py gpurun_out/${T}_launches_trainstep.csv > gpurun_out/${T}_launch_summary_trainstep.txt 2>&1; tail -1 gpurun_out/${T}_launch_summary_trainstep.txt
gzip -f gpurun_out/${T}_launches_trainstep.csv
unset VCN_BENCH_GRAPH VCN_BENCH_KERNELS
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"conv_igemm|lpips_tap|smallc|first_conv|torgb_bwd|bias_act_rows" -c 14 -o /tmp/${T}_full python tools/prof_kernels.py quick conv tconv lpips torgb first bias > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw.csv 2>/dev/null
ls -la /tmp/${T}_full.ncu-rep gpurun_out/${T}_ncu_full_raw.csv; du -sh gpurun_out
