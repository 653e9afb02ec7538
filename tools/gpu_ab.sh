#!/bin/bash
# A/B of the step with one switch flipped at a time (same box, same process order): ms per step from the bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=${1:-r2l}
export VCN_BENCH_KERNELS=0
: > gpurun_out/${T}_ab.txt
for v in "BASE=1" "VCN_VGG_FIRST=cudnn" "VCN_LPIPS_BULK=0" "VCN_BWD_PACK_SMALLC=0" "VCN_STEM_POOL=aten" "VCN_BWD_CL_BULK=0" "VCN_PREFETCH_FAKE=0" "BASE=2"; do
  env $v timeout -k 5 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ab_line.json 2> gpurun_out/${T}_ab_err.log
  python - "$v" >> gpurun_out/${T}_ab.txt <<'PY'
import json,sys
try:
    d=json.load(open('gpurun_out/'+__import__('os').environ.get('T','r2l')+'_ab_line.json'))
    print("%-24s %.4f ms/step  %.3f steps/s  launches %d" % (sys.argv[1], d['ms_per_step'], d['value'], d['gpu_launches']))
except Exception as e:
    print("%-24s FAILED %s" % (sys.argv[1], e))
PY
done
cat gpurun_out/${T}_ab.txt
timeout -k 5 600 python -m pytest tests/test_gpu_step.py tests/test_gpu_frozen.py tests/test_gpu_lpips_tap.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest_step.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_step.log
unset VCN_BENCH_KERNELS
timeout -k 5 400 python tools/run_rooflines.py > gpurun_out/${T}_rooflines.txt 2>&1
cut -c1-75,112-150 gpurun_out/${T}_rooflines.txt
