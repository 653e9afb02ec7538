#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --tb=short --maxfail=25 -p no:cacheprovider -s > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
grep -E "passed|failed|FAILED|FusedResNet|FusedVGG|graphed vs eager" gpurun_out/r2b_pytest.log | tail -30
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2b_bench_trainstep.json 2> gpurun_out/r2b_bench_trainstep.err
echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2b_bench_trainstep.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','step_tflops')}, d['e2e']['value'])"
tail -3 gpurun_out/r2b_bench_trainstep.err
VCN_VGG_IMPL=vcn VCN_BENCH_KERNELS=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench_trainstep_vggvcn.json 2> gpurun_out/r2b_bench_trainstep_vggvcn.err
python -c "
import json; d=json.load(open('gpurun_out/r2b_bench_trainstep_vggvcn.json')); print('vgg=vcn', {k:d[k] for k in ('value','ms_per_step','gpu_launches')})"
