#!/bin/bash
# round-2 pass i: conv policy experiments with the new issue loops, launch list of the step, ncu --set full of the GEMM
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r2i
timeout -k 5 600 python -m pytest tests/test_gpu_modconv.py tests/test_gpu_networks.py tests/test_zz_no_alloc.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest_conv.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_conv.log
for v in "default" "VCN_CONV_PAIR=2" "VCN_CONV_SWAP=0" "VCN_CONV_SWAP=2"; do
  echo "== $v"
  if [ "$v" = default ]; then timeout -k 5 300 python tools/run_rooflines.py conv_igemm > gpurun_out/${T}_rooflines_conv_default.txt 2>&1; cut -c1-60,112-150 gpurun_out/${T}_rooflines_conv_default.txt
  else env $v timeout -k 5 300 python tools/run_rooflines.py conv_igemm > gpurun_out/${T}_rooflines_conv_${v//=/_}.txt 2>&1; cut -c1-60,112-150 gpurun_out/${T}_rooflines_conv_${v//=/_}.txt; fi
done
export VCN_BENCH_GRAPH=0 VCN_BENCH_KERNELS=0
timeout -k 5 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain.log 2>&1 &&
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/${T}_launches_trainstep.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
echo "launchlist rc=$?"; wc -l gpurun_out/${T}_launches_trainstep.csv
python tools/launch_summary.py gpurun_out/${T}_launches_trainstep.csv > gpurun_out/${T}_launch_summary_trainstep.txt 2>&1; tail -1 gpurun_out/${T}_launch_summary_trainstep.txt
gzip -f gpurun_out/${T}_launches_trainstep.csv
unset VCN_BENCH_GRAPH VCN_BENCH_KERNELS
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -c 6 -o gpurun_out/${T}_conv_full python tools/prof_kernels.py quick conv tconv > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/${T}_conv_full.ncu-rep
