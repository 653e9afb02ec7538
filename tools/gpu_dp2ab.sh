#!/bin/bash
# 2-GPU A/B: SM reserve for the collectives x generator prefetch under the all-reduce tail
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export T=${1:-r2p}
export VCN_BENCH_KERNELS=0
: > gpurun_out/${T}_dp_ab.txt
port=29700
for v in "VCN_DP_RESERVE_SMS=8 VCN_PREFETCH_FAKE=1" "VCN_DP_RESERVE_SMS=8 VCN_PREFETCH_FAKE=0" "VCN_DP_RESERVE_SMS=0 VCN_PREFETCH_FAKE=1" "VCN_DP_RESERVE_SMS=0 VCN_PREFETCH_FAKE=0" "VCN_DP_RESERVE_SMS=16 VCN_PREFETCH_FAKE=1" "VCN_DP_RESERVE_SMS=4 VCN_PREFETCH_FAKE=1"; do
  port=$((port+1))
  env $v timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${GPUS:-2} --master-addr 127.0.0.1 --master-port $port bench.py --gpus ${GPUS:-2} --steps 30 --warmup 3 > gpurun_out/${T}_line.json 2> gpurun_out/${T}_err.log
  python - "$v" >> gpurun_out/${T}_dp_ab.txt <<'PY'
import json,sys,os
try:
    l=[x for x in open('gpurun_out/'+os.environ['T']+'_line.json') if x.startswith('{')][-1]
    d=json.loads(l)
    print("%-48s %.4f ms/step  %.3f step-GPU/s  n_gpus %d" % (sys.argv[1], d['ms_per_step'], d['value'], d['n_gpus']))
except Exception as e:
    print("%-48s FAILED %s" % (sys.argv[1], e))
PY
done
timeout -k 5 300 python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_line.json 2>/dev/null
python - "single GPU" >> gpurun_out/${T}_dp_ab.txt <<'PY'
import json,sys,os
d=json.load(open('gpurun_out/'+os.environ['T']+'_line.json'))
print("%-48s %.4f ms/step  %.3f steps/s" % (sys.argv[1], d['ms_per_step'], d['value']))
PY
cat gpurun_out/${T}_dp_ab.txt; tail -3 gpurun_out/${T}_err.log
