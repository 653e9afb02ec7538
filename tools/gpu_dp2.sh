#!/bin/bash
# 2-GPU pass: NCCL parity of the real step (tools/dp_check.py through its test) and the 2-rank bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2m}
nvidia-smi -L
timeout -k 5 600 python -m pytest tests/test_gpu_step.py -m gpu -x -q --tb=short -p no:cacheprovider -k "data_parallel" -s > gpurun_out/${T}_pytest_dp.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest_dp.log
export VCN_BENCH_KERNELS=0
for N in 2; do
  timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err
  echo "bench N=$N rc=$?"; head -c 400 gpurun_out/${T}_bench_${N}gpu.json; echo
done
VCN_PREFETCH_FAKE=0 timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/${T}_bench_2gpu_noprefetch.json 2> gpurun_out/${T}_bench_2gpu_noprefetch.err
echo "bench noprefetch rc=$?"; head -c 300 gpurun_out/${T}_bench_2gpu_noprefetch.json; echo
timeout -k 5 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_1gpu.json 2>/dev/null; head -c 300 gpurun_out/${T}_bench_1gpu.json; echo
