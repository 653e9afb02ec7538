#!/bin/bash
# forked bottleneck outputs (residual gradient adds inside the ReLU-mask kernel); new unit tests
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=r3d
timeout -k 5 900 python -m pytest tests/test_gpu_networks.py tests/test_gpu_modconv.py tests/test_gpu_step.py -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
export VCN_BENCH_KERNELS=0
: > gpurun_out/${T}_ab.txt
for v in "BASE=1" "VCN_RGB_TAP=0" "BASE=2" "VCN_RGB_TAP=0"; do
  env $v timeout -k 5 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ab_line.json 2> gpurun_out/${T}_ab_err.log
  python - "$v" >> gpurun_out/${T}_ab.txt <<'PY'
import json,sys,os
try:
    d=json.load(open('gpurun_out/'+os.environ['T']+'_ab_line.json'))
    print("%-26s %.4f ms/step  %.3f steps/s  e2e %.3f  launches %d" % (sys.argv[1], d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches']))
except Exception as e:
    print("%-26s FAILED %s" % (sys.argv[1], e))
PY
done
cat gpurun_out/${T}_ab.txt; tail -3 gpurun_out/${T}_ab_err.log
