"""Probe: does aten::cudnn_convolution_relu / _add_relu run bf16 channels_last on this stack, and how does it compare with
conv2d + bias add + relu (what autocast runs today) and with vcn conv_igemm + fused epilogue?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import viscoin_next_b200 as V
from viscoin_next_b200 import _lib
from bench import time_kernel_flushed
M = V.modconv
dev = torch.device("cuda")
torch.backends.cudnn.benchmark = False

def t(fn):
    return time_kernel_flushed(fn, dev, 10)

shapes = [  # (name, N, Cin, Cout, H, k, stride, pad)
    ("vgg1_2", 32, 64, 64, 256, 3, 1, 1), ("vgg2_2", 32, 128, 128, 128, 3, 1, 1), ("vgg3_2", 32, 256, 256, 64, 3, 1, 1),
    ("vgg4_2", 32, 512, 512, 32, 3, 1, 1), ("vgg5_2", 32, 512, 512, 16, 3, 1, 1),
    ("res2_1x1a", 16, 256, 64, 64, 1, 1, 0), ("res2_3x3", 16, 64, 64, 64, 3, 1, 1), ("res2_1x1b", 16, 64, 256, 64, 1, 1, 0),
    ("res3_3x3s2", 16, 128, 128, 64, 3, 2, 1), ("res4_3x3", 16, 256, 256, 16, 3, 1, 1), ("res5_1x1b", 16, 512, 2048, 8, 1, 1, 0),
]
for name, n, cin, cout, h, k, s, p in shapes:
    x = torch.randn(n, cin, h, h, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5).bfloat16().contiguous(memory_format=torch.channels_last)
    b = torch.randn(cout, device=dev).bfloat16()
    ref = torch.relu(F.conv2d(x.float(), w.float(), b.float(), stride=s, padding=p))
    res = {}
    res["conv+bias+relu (3 kernels)"] = t(lambda: torch.relu(F.conv2d(x, w, b, stride=s, padding=p)))
    try:
        y = torch.cudnn_convolution_relu(x, w, b, [s, s], [p, p], [1, 1], 1)
        err = float((y.float() - ref).abs().max() / ref.abs().max())
        res["cudnn_convolution_relu err %.1e cl=%s" % (err, y.is_contiguous(memory_format=torch.channels_last))] = t(
            lambda: torch.cudnn_convolution_relu(x, w, b, [s, s], [p, p], [1, 1], 1))
        z = torch.randn_like(y)
        res["cudnn_convolution_add_relu"] = t(lambda: torch.cudnn_convolution_add_relu(x, w, z, 1.0, b, [s, s], [p, p], [1, 1], 1))
    except Exception as e:
        res["cudnn fused FAILED: %s" % str(e)[:100]] = float("nan")
    if k == 3 and s == 1:
        xh = x.permute(0, 2, 3, 1)
        assert xh.is_contiguous()
        wt, _ = M.pack_weights(w.float())
        out = torch.empty(n, cout, h, h, device=dev, dtype=torch.bfloat16, memory_format=torch.channels_last)
        epi, hold = M.make_epilogue(bias=b.float(), act="relu", gain=1.0)
        M.conv_igemm(xh, wt, _lib.CONV_FPROP, cout, 3, 3, out=out, epilogue=epi)
        err = float((out.float() - ref).abs().max() / ref.abs().max())
        res["vcn conv_igemm + bias+relu epilogue err %.1e" % err] = t(lambda: M.conv_igemm(xh, wt, _lib.CONV_FPROP, cout, 3, 3, out=out, epilogue=epi))
    flop = 2.0 * n * cout * cin * k * k * (h // s) ** 2
    print("%-12s N=%d %d->%d @%d k%d s%d  (%.1f GF)" % (name, n, cin, cout, h, k, s, flop / 1e9))
    for kk, v in res.items():
        print("    %-70s %.4f ms  %.0f TF/s" % (kk, v, flop / v / 1e9 if v == v else 0))
# max pool: torch vs bytes
x = torch.randn(32, 64, 256, 256, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
ms = t(lambda: F.max_pool2d(x, 2, 2))
print("max_pool2d bf16 NHWC [32,64,256,256]: %.4f ms = %.0f GB/s algorithmic" % (ms, (x.numel() * 2 * 1.25) / ms / 1e6))
