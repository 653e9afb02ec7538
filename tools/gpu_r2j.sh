#!/bin/bash
# round-2 pass j: bulk-copy staged channels-last backward passes, swapped TCONV2, launch list, ncu --set full of the GEMM
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r2j
timeout -k 5 900 python -m pytest tests/test_gpu_modconv.py tests/test_gpu_networks.py tests/test_gpu_step.py tests/test_zz_no_alloc.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest_conv.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_conv.log
timeout -k 5 300 python tools/run_rooflines.py modconv_bwd > gpurun_out/${T}_rooflines_bwdcl.txt 2>&1; cut -c1-70,112-150 gpurun_out/${T}_rooflines_bwdcl.txt
VCN_BWD_CL_BULK=0 timeout -k 5 300 python tools/run_rooflines.py modconv_bwd > gpurun_out/${T}_rooflines_bwdcl_nobulk.txt 2>&1; cut -c1-70,112-150 gpurun_out/${T}_rooflines_bwdcl_nobulk.txt
timeout -k 5 300 python tools/run_rooflines.py TCONV2 > gpurun_out/${T}_rooflines_tconv.txt 2>&1; cut -c1-70,112-150 gpurun_out/${T}_rooflines_tconv.txt
timeout -k 5 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_trainstep.json 2> gpurun_out/${T}_bench_trainstep.err
echo "bench rc=$?"; head -c 300 gpurun_out/${T}_bench_trainstep.json; echo
export VCN_BENCH_GRAPH=0 VCN_BENCH_KERNELS=0
timeout -k 5 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain.log 2>&1 &&
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/${T}_launches_trainstep.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
echo "launchlist rc=$?"; wc -l gpurun_out/${T}_launches_trainstep.csv
python tools/launch_summary.py gpurun_out/${T}_launches_trainstep.csv > gpurun_out/${T}_launch_summary_trainstep.txt 2>&1; tail -1 gpurun_out/${T}_launch_summary_trainstep.txt
gzip -f gpurun_out/${T}_launches_trainstep.csv
unset VCN_BENCH_GRAPH VCN_BENCH_KERNELS
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -c 4 -o /tmp/${T}_conv_full python tools/prof_kernels.py quick conv tconv > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/${T}_conv_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw_conv.csv 2>/dev/null
ls -la /tmp/${T}_conv_full.ncu-rep gpurun_out/${T}_ncu_full_raw_conv.csv
du -sh gpurun_out
