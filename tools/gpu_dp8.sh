#!/bin/bash
# 8-GPU pass: the driver's scaling command at N = 8, with and without the generator prefetch
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2w}
export VCN_BENCH_KERNELS=0
N=${GPUS:-8}
date +%s.%N
timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err
echo "bench N=$N rc=$?"; date +%s.%N; grep '^{' gpurun_out/${T}_bench_${N}gpu.json | head -c 330; echo
VCN_PREFETCH_FAKE=0 timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_${N}gpu_noprefetch.json 2> gpurun_out/${T}_bench_${N}gpu_noprefetch.err
echo "noprefetch rc=$?"; date +%s.%N; grep '^{' gpurun_out/${T}_bench_${N}gpu_noprefetch.json | head -c 330; echo
tail -3 gpurun_out/${T}_bench_${N}gpu.err
