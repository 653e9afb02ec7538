"""One small call of every hand-written kernel family, for compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_ops.py [family ...]

Sizes are the smallest that still reach each kernel's fast path (the bulk-ring kernels need >= 2^21 elements).  Every
result is also compared loosely with a torch formulation so that a silent wrong answer shows up here too."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import viscoin_next_b200 as V  # noqa: E402

dev = "cuda"
what = set(sys.argv[1:])
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)
cl = lambda t: t.contiguous(memory_format=torch.channels_last)
f4 = V.upfirdn2d.setup_filter([1, 3, 3, 1], device=dev)


def on(name):
    ok = not what or name in what
    if ok:
        print("==", name, flush=True)
    return ok


def close(a, b, tol, msg):
    e = float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))
    assert e <= tol, (msg, e)
    print("   %-52s rel err %.2e" % (msg, e), flush=True)


if on("bias_act"):
    x = rn(2, 16, 40, 40).requires_grad_(True)
    b = rn(16).requires_grad_(True)
    y = V.bias_act.bias_act(x, b, act="lrelu", clamp=2.0)
    y.square().sum().backward()
    yr = torch.clamp(F.leaky_relu(x.detach() + b.detach().view(1, -1, 1, 1), 0.2) * 2 ** 0.5, -2, 2)
    close(y, yr, 1e-5, "bias_act rows fwd")
    x2 = rn(3, 7, 5)
    close(V.bias_act.bias_act(x2, rn(7) * 0, dim=1, act="tanh"), torch.tanh(x2), 1e-5, "bias_act flat fwd")

if on("upfirdn2d"):
    x = rn(2, 8, 33, 33)
    close(V.upfirdn2d.upfirdn2d(x, f4, padding=[1, 1, 1, 1], gain=4),
          F.conv2d(F.pad(x, [1, 1, 1, 1]).reshape(16, 1, 35, 35), (f4 * 4).flip([0, 1]).view(1, 1, 4, 4)).reshape(2, 8, 32, 32), 1e-5, "tile A")
    V.upfirdn2d.upfirdn2d(rn(2, 8, 32, 32), f4, padding=[2, 2, 2, 2], gain=4, flip_filter=True)
    V.upfirdn2d.upfirdn2d(rn(2, 4, 132, 132), f4, padding=[2, 2, 2, 2], gain=4, flip_filter=True)  # balanced register-staged tiles
    V.upfirdn2d.upfirdn2d(rn(2, 3, 16, 16), f4, up=2, padding=[2, 1, 2, 1], gain=4)
    V.upfirdn2d.upfirdn2d(rn(2, 8, 32, 32), f4, down=2, padding=[1, 1, 1, 1])
    xc = cl(rn(2, 128, 33, 33).bfloat16())
    close(V.upfirdn2d.upfirdn2d(xc, f4, padding=[1, 1, 1, 1], gain=4), V.upfirdn2d.upfirdn2d(xc.float(), f4, padding=[1, 1, 1, 1], gain=4), 2e-2, "cl_sep bf16")
    sym = rn(12)
    V.upfirdn2d.upfirdn2d(rn(1, 3, 64, 64), sym.unsqueeze(0), up=[2, 1], padding=[6, 5, 0, 0])
    V.upfirdn2d.upfirdn2d(rn(1, 3, 64, 64), sym.unsqueeze(1), down=[1, 2], padding=[0, 0, 6, 5])
    V.upfirdn2d.upfirdn2d(rn(1, 2, 20, 20), rn(5, 3), up=3, down=2, padding=[2, 1, 0, 3])

if on("modconv"):
    from oracle import ref_ops as R
    for (n, cin, cout, r, up) in [(2, 64, 64, 16, 1), (2, 64, 32, 16, 2), (2, 128, 128, 16, 1), (1, 16, 3, 16, 1)]:
        k = 1 if cout == 3 else 3
        x = rn(n, cin, r, r).requires_grad_(True)
        w = rn(cout, cin, k, k).requires_grad_(True)
        s = (rn(n, cin) + 1).requires_grad_(True)
        noise = rn(n, 1, r * up, r * up) * 0.1
        y = V.modconv.modulated_conv2d(x, w, s, noise=noise, up=up, padding=k // 2, resample_filter=f4 if up == 2 else None,
                                       demodulate=cout != 3)
        yr = R.modulated_conv2d(x.detach().cpu(), w.detach().cpu(), s.detach().cpu(), noise=noise.cpu(), up=up, padding=k // 2,
                                resample_filter=f4.cpu() if up == 2 else None, demodulate=cout != 3, fused_modconv=False)
        close(y, yr.to(dev), 5e-3, "modulated_conv2d %d->%d up%d (tf32)" % (cin, cout, up))
        y.square().mean().backward()

if on("resident"):
    # bf16 channels-last pipeline: bulk-ring backward passes, fused ToRGB backward, swapped / paired GEMM tiles
    for (n, cin, cout, r, up) in [(4, 128, 128, 64, 1), (4, 256, 128, 32, 2), (4, 256, 256, 64, 1)]:
        x = cl(rn(n, cin, r, r).bfloat16()).requires_grad_(True)
        w = rn(cout, cin, 3, 3)
        s = (rn(n, cin) + 1).requires_grad_(True)
        b = rn(cout)
        y = V.modconv.modconv_layer(x, w, s, bias=b, up=up, resample_filter=f4 if up == 2 else None, act="lrelu", clamp=256.0)
        assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)
        y.float().square().mean().backward()
        assert torch.isfinite(x.grad.float()).all() and torch.isfinite(s.grad).all()
    x = cl(rn(4, 128, 64, 64).bfloat16()).requires_grad_(True)
    s = (rn(4, 128) + 1).requires_grad_(True)
    w3 = rn(3, 128, 1, 1)
    b3 = rn(3)
    rgb = V.modconv.modconv_layer(x, w3, s, bias=b3, demodulate=False, act="linear", clamp=256.0)
    rgb.float().square().mean().backward()
    ref = (x.detach().float() * s.detach().view(4, 128, 1, 1))
    close(rgb, torch.clamp(F.conv2d(ref, w3, b3), -256, 256), 2e-2, "ToRGB forward (bf16 operands)")
    xr = x.detach().float().requires_grad_(True)
    sr = s.detach().clone().requires_grad_(True)
    torch.clamp(F.conv2d(xr * sr.view(4, 128, 1, 1), w3, b3), -256, 256).square().mean().backward()
    close(x.grad, xr.grad, 3e-2, "ToRGB fused backward dx")
    close(s.grad, sr.grad, 3e-2, "ToRGB fused backward dstyles")
    assert torch.isfinite(x.grad.float()).all() and torch.isfinite(s.grad).all()

if on("lpips"):
    for c, r in [(64, 96), (512, 32), (128, 16)]:
        fa = cl(rn(4, c, r, r).bfloat16()).requires_grad_(True)
        fb = cl(rn(4, c, r, r).bfloat16())
        lin = torch.rand(c, device=dev) / c
        v = V.lpips_tap.lpips_tap(fa, fb, lin)
        close(v, V.lpips_tap.lpips_tap_ref(fa.detach().float(), fb.float(), lin).reshape(-1), 2e-2, "lpips tap c=%d" % c)
        v.sum().backward()

if on("frozen"):
    import viscoin_next_b200.host.frozen_nets as fz
    x = cl(torch.relu(rn(2, 64, 32, 48)).bfloat16())
    assert torch.equal(fz.maxpool2x2(x), F.max_pool2d(x, 2))
    xa = x.clone().requires_grad_(True)
    fz.maxpool3x3s2(xa).float().sum().backward()
    gp = cl(rn(2, 64, 16, 24).bfloat16())
    fz.maxpool2x2_relu_bwd(x, gp, x)
    fz.relu_mask_bwd(x, x, x)
    img = cl(rn(2, 3, 40, 56).bfloat16())
    for ks, st, pd in [(3, 1, 1), (7, 2, 3)]:
        w = cl((rn(64, 3, ks, ks) / ks).bfloat16())
        bb = rn(64).bfloat16()
        close(fz.conv_first3(img, w, bb, ks, st, pd, True), torch.relu(F.conv2d(img.float(), w.float(), bb.float(), st, pd)), 1e-2, "conv_first3 %dx%d" % (ks, ks))

if on("wgrad"):
    x = rn(2, 64, 16, 16).requires_grad_(True)
    w = rn(64, 64, 3, 3).requires_grad_(True)
    s = (rn(2, 64) + 1).requires_grad_(True)
    V.modconv.modulated_conv2d(x, w, s, padding=1).square().mean().backward()
    assert torch.isfinite(w.grad).all()

torch.cuda.synchronize()
print("sanitize_ops: all families ran", flush=True)
