"""Timing helper: channels-last bf16 FIR (pattern A) at the synthesis sizes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import viscoin_next_b200 as V
f4 = V.upfirdn2d.setup_filter([1, 3, 3, 1], device="cuda")
for n, c, r in [(8, 128, 257), (8, 256, 129), (8, 512, 65)]:
    x = torch.randn(n, c, r, r, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    y = V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        y = V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    gb = (x.numel() + y.numel()) * 2 / 1e9
    print((n, c, r), "ms %.4f  GB/s %.0f" % (ms, gb / ms * 1e3))
