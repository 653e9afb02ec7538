"""Profiling driver: the dominant kernels at their BASELINE shapes, a few launches each (for `ncu --set full -k regex:...`).

    python tools/prof_kernels.py [conv] [tconv] [fir] [fircl] [bias]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import viscoin_next_b200 as V  # noqa: E402
from viscoin_next_b200 import _lib  # noqa: E402

M = V.modconv
what = set(a for a in sys.argv[1:] if a != "quick") or {"conv", "tconv", "fir", "fircl", "bias"}
quick = "quick" in sys.argv[1:]  # first shape of each family, one timed launch (for ncu --set full)
iters = 1 if quick else 3
dev = "cuda"
f4 = V.upfirdn2d.setup_filter([1, 3, 3, 1], device=dev)


PROF = os.environ.get("VCN_CONV_PROFILE") == "1"  # per-role cycle counters of conv_igemm (see include/vcn_b200.h)


def conv_counters(reset):
    import ctypes
    out = (ctypes.c_uint64 * 8)()
    _lib.load().vcn_conv_profile_counters(out, int(reset))
    return list(out)


GRAPH = os.environ.get("VCN_PROF_GRAPH", "1") != "0" and not quick  # replay the launches from one CUDA graph: device time only


def timed(name, fn, work, unit):
    global iters
    fn()
    torch.cuda.synchronize()
    if PROF:
        conv_counters(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if GRAPH and not PROF:
        # python + ctypes + cuTensorMapEncode cost 20-40 us per call - as much as the short kernels themselves; a graph of
        # `reps` back-to-back launches takes the host out of the measurement (the ABI is capture-safe)
        reps = 10
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for _ in range(reps):
                    fn()
        torch.cuda.current_stream().wait_stream(side)
        g.replay()
        torch.cuda.synchronize()
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
    else:
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
    print("%-46s %.4f ms  %.1f %s" % (name, ms, work / ms / 1e9 * (1e3 if unit == "GB/s" else 1), unit), flush=True)
    if PROF and name.startswith("conv"):
        c = conv_counters(True)
        t = max(c[4], 1)
        print("    per tile [cycles]: epilogue busy %.0f, epilogue waits %.0f | MMA waits: accumulator %.0f, operands %.0f | "
              "producer waits %.0f | tiles/launch %.0f, k-blocks/tile %.1f" %
              (c[0] / t, c[1] / t, c[2] / t, c[3] / t, c[5] / t, t / iters, c[6] / t), flush=True)


if "conv" in what:
    for n, cin, cout, r in [(8, 128, 128, 256), (8, 256, 256, 128), (8, 512, 512, 64)][:1 if quick else 3]:
        x = torch.randn(n, cin, r, r, device=dev)
        w = torch.randn(cout, cin, 3, 3, device=dev)
        xh = M.modulate_pack(x, None)
        wt, _ = M.pack_weights(w)
        for dt, cl in [(torch.float32, False), (torch.bfloat16, True)]:
            y = M.conv_igemm(xh, wt, _lib.CONV_FPROP, cout, 3, 3, out_dtype=dt, channels_last=cl)
            timed("conv FPROP %d->%d @%d %s" % (cin, cout, r, "bf16 NHWC" if cl else "f32 NCHW"),
                  lambda: M.conv_igemm(xh, wt, _lib.CONV_FPROP, cout, 3, 3, out=y), 2.0 * n * cout * cin * 9 * r * r, "TFLOP/s")
        del x, w, xh, wt, y

if "tconv" in what:
    for n, cin, cout, r in [(8, 256, 128, 128), (8, 512, 256, 64)][:1 if quick else 2]:
        x = torch.randn(n, cin, r, r, device=dev)
        w = torch.randn(cout, cin, 3, 3, device=dev)
        xh = M.modulate_pack(x, None)
        wt, _ = M.pack_weights(w)
        for dt, cl in [(torch.float32, False), (torch.bfloat16, True)]:
            y = M.conv_igemm(xh, wt, _lib.CONV_TCONV2, cout, 3, 3, out_dtype=dt, channels_last=cl)
            timed("conv TCONV2 %d->%d @%d %s" % (cin, cout, r, "bf16 NHWC" if cl else "f32 NCHW"),
                  lambda: M.conv_igemm(xh, wt, _lib.CONV_TCONV2, cout, 3, 3, out=y), 2.0 * n * cout * cin * 9 * r * r, "TFLOP/s")
        del x, w, xh, wt, y

if "fir" in what:
    for n, c, r in [(8, 128, 257), (8, 256, 129), (8, 512, 65)][:1 if quick else 3]:
        x = torch.randn(n, c, r, r, device=dev)
        y = V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4)
        timed("upfirdn2d A f32 NCHW [%d,%d,%d,%d]" % (n, c, r, r),
              lambda: V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4), (x.numel() + y.numel()) * 4, "GB/s")
        del x, y
    x = torch.randn(8, 128, 256, 256, device=dev)
    y = V.upfirdn2d.upfirdn2d(x, f4, padding=[2, 2, 2, 2], gain=4, flip_filter=True)
    timed("upfirdn2d A' f32 NCHW [8,128,256,256]", lambda: V.upfirdn2d.upfirdn2d(x, f4, padding=[2, 2, 2, 2], gain=4, flip_filter=True),
          (x.numel() + y.numel()) * 4, "GB/s")
    del x, y

if "fircl" in what:
    for n, c, r in [(8, 128, 257), (8, 256, 129), (8, 512, 65)][:1 if quick else 3]:
        x = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        y = V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4)
        timed("upfirdn2d A bf16 NHWC [%d,%d,%d,%d]" % (n, c, r, r),
              lambda: V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4), (x.numel() + y.numel()) * 2, "GB/s")
        del x, y

if "bias" in what:
    x = torch.randn(8, 128, 256, 256, device=dev)
    b = torch.randn(128, device=dev)
    timed("bias_act lrelu fwd f32 [8,128,256,256]", lambda: V.bias_act.bias_act(x, b, act="lrelu"), 2 * x.numel() * 4, "GB/s")

if "bwdcl" in what:
    for n, c, r in [(16, 512, 16), (16, 512, 32)]:  # low-resolution layers: latency-bound, few CTAs
        y = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        dy = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        d = torch.rand(n, c, device=dev) + 0.5
        b = torch.randn(c, device=dev)
        epi, hold = M.make_epilogue(scale=d, bias=b, act="lrelu")
        timed("bwd_pack_cl bf16 [%d,%d,%d,%d]" % (n, c, r, r), lambda: M.bwd_pack(dy, y, epi, True, True), 3 * y.numel() * 2, "GB/s")
        dxh = dy.clone()
        timed("bwd_finish_cl bf16 [%d,%d,%d,%d]" % (n, c, r, r), lambda: M.bwd_finish_cl(dxh, y, d, True), 3 * y.numel() * 2, "GB/s")
    n, c, r = 16, 128, 256
    y = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    d = torch.rand(n, c, device=dev) + 0.5
    b = torch.randn(c, device=dev)
    epi, hold = M.make_epilogue(scale=d, bias=b, act="lrelu", clamp=256.0)
    timed("bwd_pack_cl bf16 [16,128,256,256]", lambda: M.bwd_pack(dy, y, epi, True, True), 3 * y.numel() * 2, "GB/s")
    s = torch.rand(n, c, device=dev) + 0.5
    dxh = dy.clone()
    timed("bwd_finish_cl bf16 [16,128,256,256]", lambda: M.bwd_finish_cl(dxh, y, s, True), 3 * y.numel() * 2, "GB/s")
    x32 = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    timed("modulate_pack_cl bf16 [16,128,256,256]", lambda: M.modulate_pack(x32, s), 2 * y.numel() * 2, "GB/s")

if "lpips" in what:
    n, c, r = 16, 64, 256
    fa = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    fb = torch.randn(n, c, r, r, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    lin = torch.rand(c, device=dev) / c
    timed("lpips_tap fwd bf16 [16,64,256,256]", lambda: V.lpips_tap.lpips_tap(fa.detach(), fb, lin), 2 * fa.numel() * 2, "GB/s")
    iters = 1
    GRAPH = False
    timed("lpips_tap fwd+bwd bf16 [16,64,256,256]", lambda: V.lpips_tap.lpips_tap(fa, fb, lin).sum().backward(), 5 * fa.numel() * 2, "GB/s")

if "torgb" in what:
    n = 16
    dimg, yimg = torch.randn(n, 3, 256, 256, device=dev), torch.randn(n, 3, 256, 256, device=dev)
    epi3, hold3 = M.make_epilogue(bias=torch.randn(3, device=dev), act="linear", clamp=256.0)
    timed("bwd_pack small-c ToRGB f32 [16,3,256,256]", lambda: M.bwd_pack(dimg, yimg, epi3, False, True), 2 * dimg.numel() * 4 + n * 65536 * 16, "GB/s")

    xr = torch.randn(n, 128, 256, 256, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    dy0r = torch.randn(n, 256, 256, 8, device=dev).bfloat16()
    w3 = torch.randn(3, 128, 1, 1, device=dev)
    s3 = torch.rand(n, 128, device=dev) + 0.5
    timed("torgb_bwd_fused bf16 [16,128,256,256]", lambda: M.torgb_bwd_fused(xr, dy0r, w3, s3, True), 2 * xr.numel() * 2 + dy0r.numel() * 2, "GB/s")

if "first" in what:
    import viscoin_next_b200.host.frozen_nets as fz
    img = torch.randn(16, 3, 256, 256, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    for ks, st_, pd in [(3, 1, 1), (7, 2, 3)]:
        w = (torch.randn(64, 3, ks, ks, device=dev) / ks).bfloat16().contiguous(memory_format=torch.channels_last)
        bb = torch.randn(64, device=dev).bfloat16()
        oh = (256 + 2 * pd - ks) // st_ + 1
        timed("conv_first3 %dx%d/s%d bf16 [16,3,256,256]->64ch" % (ks, ks, st_), lambda: fz.conv_first3(img, w, bb, ks, st_, pd, True),
              16 * oh * oh * 64 * 2 + img.numel() * 2, "GB/s")
