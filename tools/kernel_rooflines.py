"""Per-kernel roofline rows for bench.py (`roofline_others`): every hand-written kernel of the hot path alone at its
BASELINE shape, L2 flushed before every launch (bench.time_kernel_flushed), against MEASURED_PEAKS.json - and, for the
convolutions, the cuDNN kernel the reference's GPU path runs for the same layer (F.conv2d / F.conv_transpose2d, bf16
channels_last and fp32-TF32 NCHW, same timing method): SURVEY 8(d) "the kernel to beat on the same box".

Algorithmic work per launch: bias_act fwd 2*S*e, grad-1 3*S*e; upfirdn2d (in + out)*e; conv 2*N*Cout*Cin*k^2*H*W
(transposed: per INPUT pixel); bwd_pack / bwd_finish 3*S*e; modulate_pack 2*S*e."""

import json
import os

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _traffic():
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            return json.load(fh)
    except Exception:
        return {}


def collect(device, peaks, time_fn, iters=10, include_cudnn=True, n=8):
    import viscoin_next_b200 as V
    from viscoin_next_b200 import _lib
    M = V.modconv
    rows = []
    traffic = _traffic()
    f4 = V.upfirdn2d.setup_filter([1, 3, 3, 1], device=device)

    def hbm(kernel, fn, nbytes, key=None, **extra):
        ms = time_fn(fn, device, iters)
        ach = nbytes / (ms * 1e-3) / 1e9
        rows.append(dict(kernel=kernel, bound="hbm", achieved=round(ach, 1), peak=peaks["hbm"], unit="GB/s",
                         frac=round(ach / peaks["hbm"], 4), alg_bytes_per_launch=int(nbytes), avg_launch_ms=round(ms, 5),
                         traffic=traffic.get(key) if key else None, **extra))

    def tensor(kernel, fn, flop, key=None, **extra):
        ms = time_fn(fn, device, iters)
        ach = flop / (ms * 1e-3) / 1e12
        row = dict(kernel=kernel, bound="tensor", achieved=round(ach, 1), peak=peaks["bf16"], unit="TFLOP/s",
                   frac=round(ach / peaks["bf16"], 4), alg_flop_per_launch=flop, avg_launch_ms=round(ms, 5),
                   traffic=traffic.get(key) if key else None)
        row.update(extra)
        rows.append(row)
        return row

    g = torch.Generator(device=device).manual_seed(0)
    rn = lambda *s: torch.randn(*s, device=device, generator=g)

    # ---- bias_act (fp32 NCHW, the reference's layout) ---------------------------------------------------------------
    x = rn(n, 128, 256, 256)
    b = rn(128)
    hbm("bias_act_rows_kernel<float,lrelu> fwd [%d,128,256,256] fp32 NCHW" % n, lambda: V.bias_act.bias_act(x, b, act="lrelu"),
        2 * x.numel() * 4, key="bias_act_rows_kernel")
    y = V.bias_act.bias_act(x, b, act="lrelu")
    dy = rn(n, 128, 256, 256)
    spec = V.bias_act.activation_funcs["lrelu"]
    hbm("bias_act_rows_kernel<float,lrelu> grad-1 + fused db [%d,128,256,256] fp32 NCHW" % n,
        lambda: V.bias_act._call(dy, None, None, y, None, 1, 1, spec.cuda_idx, spec.def_alpha, spec.def_gain, -1.0, True),
        3 * x.numel() * 4, key="bias_act_rows_kernel_grad1")
    del y, dy

    # ---- upfirdn2d patterns of SURVEY 8(a6) -------------------------------------------------------------------------
    xa = rn(n, 128, 257, 257)
    hbm("upfirdn2d_tile_kernel<float> pattern A [%d,128,257,257]->256^2 fp32 NCHW" % n,
        lambda: V.upfirdn2d.upfirdn2d(xa, f4, padding=[1, 1, 1, 1], gain=4), (xa.numel() + n * 128 * 256 * 256) * 4,
        key="upfirdn2d_tile_kernel")
    hbm("upfirdn2d pattern A' (backward of A) [%d,128,256,256]->257^2 fp32 NCHW" % n,
        lambda: V.upfirdn2d.upfirdn2d(x, f4, padding=[2, 2, 2, 2], gain=4, flip_filter=True), (x.numel() + xa.numel()) * 4,
        key="upfirdn2d_pattern_a_prime")
    xb = rn(n, 3, 128, 128)
    hbm("upfirdn2d pattern B (RGB skip, up 2) [%d,3,128,128]->256^2 fp32 NCHW (L2-resident: reported, not gated)" % n,
        lambda: V.upfirdn2d.upfirdn2d(xb, f4, up=2, padding=[2, 1, 2, 1], gain=4), (xb.numel() + n * 3 * 256 * 256) * 4)
    hbm("upfirdn2d pattern B' (down 2) [%d,128,256,256]->128^2 fp32 NCHW" % n,
        lambda: V.upfirdn2d.upfirdn2d(x, f4, down=2, padding=[1, 1, 1, 1], gain=4, flip_filter=True),
        (x.numel() + n * 128 * 128 * 128) * 4, key="upfirdn2d_pattern_b_prime")
    xcl = xa.bfloat16().contiguous(memory_format=torch.channels_last)
    hbm("upfirdn2d_cl_sep_kernel<bf16> pattern A [%d,128,257,257] bf16 NHWC" % n,
        lambda: V.upfirdn2d.upfirdn2d(xcl, f4, padding=[1, 1, 1, 1], gain=4), (xcl.numel() + n * 128 * 256 * 256) * 2,
        key="upfirdn2d_cl_sep_kernel")
    # separable long filter (AugmentPipe's 12-tap sym6 at twice the image resolution): one 1-D pass each
    sym6 = torch.tensor([0.0154, 0.0035, -0.118, -0.0483, 0.4911, 0.7876, 0.3379, -0.0726, -0.0211, 0.0447, 0.0018, -0.0078], device=device)
    xs_ = rn(n, 3, 1024, 1024)
    hbm("upfirdn1d_x_kernel<float,12 taps> [%d,3,1024,1024] fp32 NCHW (one separable pass along x)" % n,
        lambda: V.upfirdn2d.launch(xs_, sym6.unsqueeze(0), 1, 1, 1, 1, 6, 5, 0, 0, False, 1.0), 2 * xs_.numel() * 4)
    hbm("upfirdn1d_y_kernel<float,12 taps> [%d,3,1024,1024] fp32 NCHW (one separable pass along y)" % n,
        lambda: V.upfirdn2d.launch(xs_, sym6.unsqueeze(1), 1, 1, 1, 1, 0, 0, 6, 5, False, 1.0), 2 * xs_.numel() * 4)
    xh_ = xs_[:, :, :, :512].contiguous()
    hbm("upfirdn1d_x_kernel<float,12 taps, up 2> [%d,3,1024,512]->[.,.,1024,1024] fp32 NCHW" % n,
        lambda: V.upfirdn2d.launch(xh_, sym6.unsqueeze(0), 2, 1, 1, 1, 6, 5, 0, 0, False, 2.0), (xh_.numel() + xs_.numel()) * 4)
    del xa, xb, xcl, x, xs_, xh_

    # ---- channels-last passes around the GEMMs (bf16, the train step's N = 2*batch) ----------------------------------
    n2 = 2 * n
    ycl = rn(n2, 128, 256, 256).bfloat16().contiguous(memory_format=torch.channels_last)
    dycl = rn(n2, 128, 256, 256).bfloat16().contiguous(memory_format=torch.channels_last)
    d = torch.rand(n2, 128, device=device) + 0.5
    bb = rn(128)
    epi, hold = M.make_epilogue(scale=d, bias=bb, act="lrelu", clamp=256.0)
    hbm("modconv_bwd_cl_kernel<bf16,pack> [%d,128,256,256] bf16 NHWC" % n2, lambda: M.bwd_pack(dycl, ycl, epi, True, True),
        3 * ycl.numel() * 2, key="modconv_bwd_cl_pack")
    dxh = dycl.clone()
    hbm("modconv_bwd_cl_kernel<bf16,finish> [%d,128,256,256] bf16 NHWC" % n2, lambda: M.bwd_finish_cl(dxh, ycl, d, True),
        3 * ycl.numel() * 2, key="modconv_bwd_cl_finish")
    hbm("modulate_pack_cl_rows_kernel<bf16> [%d,128,256,256] bf16 NHWC" % n2, lambda: M.modulate_pack(ycl, d), 2 * ycl.numel() * 2,
        key="modulate_pack_cl_rows_kernel")
    del dxh, hold
    # LPIPS tap (relu1_2 of VGG-16: 64 channels at full resolution), forward and backward, and the ToRGB backward pack
    lin = torch.rand(64, device=device) / 64
    fa = ycl[:, :64].contiguous(memory_format=torch.channels_last).requires_grad_(True)
    fb = dycl[:, :64].contiguous(memory_format=torch.channels_last)
    hbm("lpips_tap_kernel<bf16,fwd> [%d,64,256,256] x2 bf16 NHWC" % n2, lambda: V.lpips_tap.lpips_tap(fa.detach(), fb, lin),
        2 * fa.numel() * 2, key="lpips_tap_fwd")
    import ctypes
    gv = torch.ones(n2, device=device)
    da = torch.empty_like(fb)
    fad = fa.detach()
    s4 = lambda t: ctypes.byref((ctypes.c_int64 * 4)(*[int(v) for v in t.stride()]))
    lib = _lib.load()

    def lp_bwd():
        _lib.check(lib.vcn_lpips_tap_bwd(_lib.ptr(fad), _lib.ptr(fb), _lib.dtype_code(fad.dtype), s4(fad), s4(fb), _lib.ptr(lin),
                                         _lib.ptr(gv), _lib.ptr(da), n2, 64, 256, 256, _lib.current_stream(fad.device)), "lpips_tap_bwd")

    hbm("lpips_tap_kernel<bf16,bwd> [%d,64,256,256] bf16 NHWC (2 reads + 1 write)" % n2, lp_bwd, 3 * fa.numel() * 2, key="lpips_tap_bwd")
    del fa, fb, fad, da
    dimg, yimg = rn(n2, 3, 256, 256), rn(n2, 3, 256, 256)
    epi3, hold3 = M.make_epilogue(bias=rn(3), act="linear", clamp=256.0)
    hbm("modconv_bwd_pack_smallc_kernel<float,bf16> ToRGB [%d,3,256,256] fp32 NCHW -> bf16 NHWC8" % n2,
        lambda: M.bwd_pack(dimg, yimg, epi3, False, True), (2 * dimg.numel() * 4 + n2 * 65536 * 16), key="modconv_bwd_pack_smallc")
    xr = ycl
    dy0r = rn(n2, 256, 256, 8).bfloat16()
    w3 = rn(3, 128, 1, 1)
    hbm("torgb_bwd_bulk_kernel<bf16> fused ToRGB backward [%d,128,256,256] bf16 NHWC (read x, write dx)" % n2,
        lambda: M.torgb_bwd_fused(xr, dy0r, w3, d, True), 2 * xr.numel() * 2 + dy0r.numel() * 2, key="torgb_bwd_fused")
    import viscoin_next_b200.host.frozen_nets as fz
    img = rn(n2, 3, 256, 256).bfloat16().contiguous(memory_format=torch.channels_last)
    for ks, st_, pd in [(3, 1, 1), (7, 2, 3)]:
        wf = (rn(64, 3, ks, ks) / ks).bfloat16().contiguous(memory_format=torch.channels_last)
        bf = rn(64).bfloat16()
        oh = (256 + 2 * pd - ks) // st_ + 1
        hbm("first_conv_kernel<%d,%d> 3->64 ch + bias + ReLU [%d,3,256,256] bf16 NHWC (write-bound)" % (ks, st_, n2),
            lambda: fz.conv_first3(img, wf, bf, ks, st_, pd, True), n2 * oh * oh * 64 * 2 + img.numel() * 2, key="first_conv_%d" % ks)
    del ycl, dycl, dimg, yimg, hold3, xr, dy0r, img

    # ---- the convolutions: vcn_conv_igemm vs cuDNN on the layer shapes of SURVEY App. A (res >= 64: 91 % of the FLOPs) ---
    def cudnn_ms(kind, xin, w):
        """The library call the reference makes for this layer (conv2d_gradfix.py:38,43), same timing method."""
        if kind == "fprop":
            fn = lambda: F.conv2d(xin, w, padding=1)
        elif kind == "tconv":
            fn = lambda: F.conv_transpose2d(xin, w.transpose(0, 1), stride=2)
        else:  # input gradient of a stride-1 3x3 conv = transposed conv, stride 1, padding 1
            fn = lambda: F.conv_transpose2d(xin, w, padding=1)
        return time_fn(fn, device, iters)

    shapes = [("fprop", n, 128, 128, 256), ("fprop", n, 256, 256, 128), ("fprop", n, 512, 512, 64),
              ("tconv", n, 256, 128, 128), ("tconv", n, 512, 256, 64), ("tconv", n, 512, 512, 32),
              ("dgrad", n2, 128, 128, 256), ("dgrad", n2, 256, 256, 128)]
    for kind, nn, cin, cout, r in shapes:
        xin = rn(nn, cin, r, r)
        w = rn(cout, cin, 3, 3)
        xh = M.modulate_pack(xin, None)
        if kind == "dgrad":  # dx = conv^T(dy): dy has `cout` channels, the GEMM contracts over them
            wt, _ = M.pack_weights(w, transpose_io=True)
            xh = M.modulate_pack(rn(nn, cout, r, r), None)
            ck, oc = _lib.CONV_DGRAD, cin
        else:
            wt, _ = M.pack_weights(w)
            ck, oc = (_lib.CONV_FPROP if kind == "fprop" else _lib.CONV_TCONV2), cout
        orr = 2 * r + 1 if kind == "tconv" else r
        out = torch.empty(nn, oc, orr, orr, device=device, dtype=torch.bfloat16, memory_format=torch.channels_last)
        bias = torch.zeros(oc, device=device)
        epi, hold = (M.make_epilogue(bias=bias, act="lrelu") if kind == "fprop" else (None, None))
        flop = 2.0 * nn * cout * cin * 9 * r * r
        extra = {}
        if include_cudnn:
            try:
                if kind == "dgrad":
                    xl = rn(nn, cout, r, r)
                else:
                    xl = xin
                xbf = xl.bfloat16().contiguous(memory_format=torch.channels_last)
                wbf = w.bfloat16().contiguous(memory_format=torch.channels_last)
                t_bf = cudnn_ms(kind, xbf, wbf)
                t_tf = cudnn_ms(kind, xl, w)  # fp32 NCHW, torch default cudnn.allow_tf32 = True: the reference's GPU arithmetic
                extra = {"cudnn_bf16_nhwc_ms": round(t_bf, 5), "cudnn_tf32_nchw_ms": round(t_tf, 5)}
                del xbf, wbf, xl
            except Exception as e:  # pragma: no cover
                extra = {"cudnn_error": str(e)[:120]}
        label = {"fprop": "FPROP (+bias+lrelu epilogue)", "tconv": "TCONV2 (stride-2 transposed, 4 phases)", "dgrad": "DGRAD"}[kind]
        row = tensor("conv_igemm_kernel %s [%d,%d,%d,%d]x[%d,%d,3,3] bf16 NHWC -> bf16 NHWC" % (label, nn, cin, r, r, cout, cin),
                     lambda: M.conv_igemm(xh, wt, ck, oc, 3, 3, out=out, epilogue=epi), flop,
                     key="conv_igemm_kernel" if (kind, cin, r) == ("fprop", 128, 256) else None, **extra)
        if "cudnn_bf16_nhwc_ms" in row:
            row["vs_cudnn_bf16"] = round(row["cudnn_bf16_nhwc_ms"] / row["avg_launch_ms"], 3)
            row["vs_cudnn_tf32"] = round(row["cudnn_tf32_nchw_ms"] / row["avg_launch_ms"], 3)
        del xin, w, xh, wt, out, hold
    return rows
