#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
echo "== forced pair mode (VCN_CONV_PAIR=2)"
VCN_CONV_PAIR=2 timeout -k 5 240 python -m pytest tests/test_gpu_modconv.py -x -q --tb=short -p no:cacheprovider > gpurun_out/r2d_pair_forced.log 2>&1
echo "rc=$?"; tail -15 gpurun_out/r2d_pair_forced.log
echo "== default policy"
timeout -k 5 300 python -m pytest tests/test_gpu_modconv.py tests/test_gpu_networks.py -x -q --tb=short -p no:cacheprovider > gpurun_out/r2d_pair_default.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/r2d_pair_default.log
echo "== kernel rooflines (default policy, then pair off)"
timeout -k 5 300 python tools/run_rooflines.py > gpurun_out/r2d_rooflines_pair.txt 2>&1; tail -12 gpurun_out/r2d_rooflines_pair.txt
VCN_CONV_PAIR=0 timeout -k 5 300 python tools/run_rooflines.py conv > gpurun_out/r2d_rooflines_nopair.txt 2>&1; tail -9 gpurun_out/r2d_rooflines_nopair.txt
VCN_CONV_PAIR=2 timeout -k 5 300 python tools/run_rooflines.py conv > gpurun_out/r2d_rooflines_pairall.txt 2>&1; tail -9 gpurun_out/r2d_rooflines_pairall.txt
