"""Print the per-kernel roofline rows of bench.py (tools/kernel_rooflines.collect) without running the train step."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from tools import kernel_rooflines
peaks = bench.measured_peaks()
rows = kernel_rooflines.collect(torch.device("cuda"), peaks, bench.time_kernel_flushed, iters=10, n=8)
only = sys.argv[1] if len(sys.argv) > 1 else ""
for r in rows:
    if only and only not in r["kernel"]:
        continue
    extra = {k: v for k, v in r.items() if "cudnn" in k}
    print("%-112s %8.1f %-7s frac %.3f  %.4f ms %s" % (r["kernel"][:112], r["achieved"], r["unit"], r["frac"], r["avg_launch_ms"], extra))
