#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -m gpu -q --tb=short --maxfail=25 -p no:cacheprovider -s > gpurun_out/r2f_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
grep -E "passed|failed|FAILED|operands\] worst" gpurun_out/r2f_pytest.log | tail -14
timeout -k 5 200 python tools/run_rooflines.py upfirdn1d 2>&1 | tail -4
