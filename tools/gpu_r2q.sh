#!/bin/bash
# round-2 pass q: fused ToRGB backward; full GPU suite; A/B in the step
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=r2q
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/${T}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -E "passed|failed|FAILED|Error|rc=" gpurun_out/${T}_pytest.log | tail -6
export VCN_BENCH_KERNELS=0
: > gpurun_out/${T}_ab.txt
for v in "BASE=1" "VCN_TORGB_FUSED=0" "BASE=2" "VCN_TORGB_FUSED=0"; do
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
