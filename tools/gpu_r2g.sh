#!/bin/bash
# round-2 pass g: the allocation-free entry point alone, the whole GPU suite, then the bench lines of both arms
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/r2g_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r2g_noalloc_alone.log
tail -3 gpurun_out/r2g_noalloc_alone.log
timeout -k 5 1500 python -m pytest tests -m gpu -x -q --tb=short -p no:cacheprovider > gpurun_out/r2g_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
grep -E "passed|failed|FAILED|rc=" gpurun_out/r2g_pytest.log | tail -8
timeout -k 5 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2g_bench_trainstep.json 2> gpurun_out/r2g_bench_trainstep.err
echo "bench rc=$?"; head -c 600 gpurun_out/r2g_bench_trainstep.json; tail -3 gpurun_out/r2g_bench_trainstep.err
timeout -k 5 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2g_bench_reference_arm.json 2> gpurun_out/r2g_bench_reference_arm.err
echo "ref rc=$?"; head -c 400 gpurun_out/r2g_bench_reference_arm.json
