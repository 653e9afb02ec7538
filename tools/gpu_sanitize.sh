#!/bin/bash
# compute-sanitizer over one small call of every hand-written kernel family (tools/sanitize_ops.py)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=${1:-r2s}
timeout -k 5 300 python tools/sanitize_ops.py > gpurun_out/${T}_sanitize_plain.log 2>&1; echo "plain rc=$?"; tail -4 gpurun_out/${T}_sanitize_plain.log
for tool in memcheck synccheck racecheck; do
  timeout -k 5 1200 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_ops.py > gpurun_out/${T}_sanitize_${tool}.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|all families ran|Error|hazard" gpurun_out/${T}_sanitize_${tool}.log | sort | uniq -c | head -12
done
VCN_CONV_PAIR=2 timeout -k 5 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_ops.py resident modconv > gpurun_out/${T}_sanitize_memcheck_pair.log 2>&1
echo "memcheck pair rc=$?"; grep -E "ERROR SUMMARY|all families ran" gpurun_out/${T}_sanitize_memcheck_pair.log
for f in gpurun_out/${T}_sanitize_*.log; do head -c 200000 $f > $f.tmp && mv $f.tmp $f; done
