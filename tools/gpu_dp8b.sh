#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2y}
export VCN_BENCH_KERNELS=0
N=8
timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29821 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err
echo "bench bf16-comm N=$N rc=$?"; grep '^{' gpurun_out/${T}_bench_${N}gpu.json | head -c 300; echo
VCN_DP_BF16_COMM=0 timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29822 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_${N}gpu_fp32comm.json 2> gpurun_out/${T}_bench_${N}gpu_fp32comm.err
echo "bench fp32-comm rc=$?"; grep '^{' gpurun_out/${T}_bench_${N}gpu_fp32comm.json | head -c 300; echo
