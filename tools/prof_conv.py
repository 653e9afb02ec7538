"""Profiling helper: run vcn_conv_igemm on one layer shape a few times (for ncu / CUDA-event timing)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from viscoin_next_b200 import _lib
from viscoin_next_b200.host import modconv as M

n, cin, cout, res, k, kind = [int(v) for v in sys.argv[1:7]]
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 5
x = torch.randn(n, cin, res, res, device="cuda")
w = torch.randn(cout, cin, k, k, device="cuda")
xh = M.modulate_pack(x, None)
wt, _ = M.pack_weights(w)
y = M.conv_igemm(xh, wt, kind, cout, k, k)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    M.conv_igemm(xh, wt, kind, cout, k, k, out=y)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
oh = y.shape[2]
flop = 2.0 * n * cout * cin * k * k * (res * res if kind != 1 else res * res)
print("shape", (n, cin, cout, res, k, kind), "ms %.4f  TFLOP/s %.1f" % (ms, flop / ms / 1e9))
