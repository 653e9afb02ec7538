#!/bin/bash
# round-2 GPU pass 1: full GPU test suite, bench lines of all arms, reference-GPU baselines
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --tb=short --maxfail=25 -p no:cacheprovider > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench_trainstep.json 2> gpurun_out/r2a_bench_trainstep.err
echo "bench rc=$?"; tail -c 600 gpurun_out/r2a_bench_trainstep.json
timeout 600 python bench.py --impl reference-gpu --steps 5 --warmup 3 > gpurun_out/r2a_bench_trainstep_refgpu.json 2> gpurun_out/r2a_bench_trainstep_refgpu.err
echo "refgpu rc=$?"; tail -c 900 gpurun_out/r2a_bench_trainstep_refgpu.json; tail -5 gpurun_out/r2a_bench_trainstep_refgpu.err
timeout 300 python bench.py --workload synthesis --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench_synthesis.json 2> gpurun_out/r2a_bench_synthesis.err
VCN_BENCH_KERNELS=0 timeout 300 python bench.py --workload synthesis --impl reference-gpu --steps 10 --warmup 3 > gpurun_out/r2a_bench_synthesis_refgpu.json 2> gpurun_out/r2a_bench_synthesis_refgpu.err
tail -c 700 gpurun_out/r2a_bench_synthesis_refgpu.json
VCN_BENCH_KERNELS=0 timeout 300 python bench.py --workload opsweep --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench_opsweep.json 2> gpurun_out/r2a_bench_opsweep.err
timeout 300 python bench.py --workload opsweep --impl reference-gpu --steps 20 --warmup 3 > gpurun_out/r2a_bench_opsweep_refgpu.json 2> gpurun_out/r2a_bench_opsweep_refgpu.err
tail -c 500 gpurun_out/r2a_bench_opsweep_refgpu.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_reference_arm.json 2> gpurun_out/r2a_bench_reference_arm.err
tail -c 600 gpurun_out/r2a_bench_reference_arm.json
