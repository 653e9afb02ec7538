#!/bin/bash
# round-2 pass o: direct 3-channel first convolutions, templated LPIPS ring kernel; A/B in the step; ncu --set full of the
# memory-bound kernels that are still below the gate
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=r2o
timeout -k 5 900 python -m pytest tests/test_gpu_frozen.py tests/test_gpu_lpips_tap.py tests/test_gpu_step.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
export VCN_BENCH_KERNELS=0
: > gpurun_out/${T}_ab.txt
for v in "BASE=1" "VCN_VGG_FIRST=cudnn" "BASE=2"; do
  env $v timeout -k 5 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ab_line.json 2> gpurun_out/${T}_ab_err.log
  python - "$v" >> gpurun_out/${T}_ab.txt <<'PY'
import json,sys,os
try:
    d=json.load(open('gpurun_out/'+os.environ['T']+'_ab_line.json'))
    print("%-24s %.4f ms/step  %.3f steps/s  launches %d" % (sys.argv[1], d['ms_per_step'], d['value'], d['gpu_launches']))
except Exception as e:
    print("%-24s FAILED %s" % (sys.argv[1], e))
PY
done
cat gpurun_out/${T}_ab.txt
unset VCN_BENCH_KERNELS
timeout -k 5 300 python tools/run_rooflines.py lpips > gpurun_out/${T}_rooflines_lpips.txt 2>&1; cut -c1-75,112-150 gpurun_out/${T}_rooflines_lpips.txt
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"lpips_tap|upfirdn2d_cl_sep|smallc|first_conv|modconv_bwd_cl_bulk" -c 8 -o /tmp/${T}_mem_full python tools/prof_kernels.py quick lpips fircl torgb bwdcl > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/${T}_mem_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw_mem.csv 2>/dev/null
ls -la /tmp/${T}_mem_full.ncu-rep gpurun_out/${T}_ncu_full_raw_mem.csv; du -sh gpurun_out
