"""Oracle: functional restatement of the networks that consume the hot-path ops.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  The networks are driven
by a plain ``state_dict`` with the reference's own key names, so the very same
tensors can be fed to the reference modules (in ``make_golden.py``), to this
oracle and to the CUDA product path.  Citations are relative to
``/root/reference``.
"""

import math

import torch

from . import ref_ops as R

# ---------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------


def normalize_2nd_moment(x, dim=1, eps=1e-8):
    """stylegan2_ada/training/networks.py:21-22."""
    return x * (x.square().mean(dim=dim, keepdim=True) + eps).rsqrt()


def fully_connected(x, weight, bias, activation="linear", lr_multiplier=1.0):
    """FullyConnectedLayer.forward, networks.py:105-118: runtime weight scale
    lr_mult/sqrt(in), bias scale lr_mult; addmm when linear+bias, else matmul+bias_act."""
    w = weight.to(x.dtype) * (lr_multiplier / math.sqrt(weight.shape[1]))
    b = bias
    if b is not None:
        b = b.to(x.dtype)
        if lr_multiplier != 1:
            b = b * lr_multiplier
    if activation == "linear" and b is not None:
        return torch.addmm(b.unsqueeze(0), x, w.t())
    return R.bias_act(x.matmul(w.t()), b, act=activation)


def synthesis_channels(img_resolution, channel_base=32768, channel_max=512):
    """networks.py:439-441 -> {res: channels}."""
    log2 = int(math.log2(img_resolution))
    return {2 ** i: min(channel_base // (2 ** i), channel_max) for i in range(2, log2 + 1)}


def synthesis_num_ws(img_resolution):
    """networks.py:444-455: one w per conv + one for the last torgb."""
    log2 = int(math.log2(img_resolution))
    return 2 * (log2 - 1)


# ---------------------------------------------------------------------------
# synthesis side
# ---------------------------------------------------------------------------


def _noise_for(sd, p, n, res, noise_mode, noises, device):
    """SynthesisLayer.forward noise selection, networks.py:292-296."""
    if noise_mode == "none":
        return None
    strength = sd[p + "noise_strength"]
    if noise_mode == "const":
        return sd[p + "noise_const"] * strength
    if noises is not None:  # explicit per-layer noise keeps RNG out of parity tests
        return noises[p] * strength
    return torch.randn([n, 1, res, res], device=device) * strength


def synthesis_layer(x, w, sd, p, res, up, noise_mode, fused_modconv, gain=1.0, conv_clamp=None,
                    noises=None, activation="lrelu", use_noise=True):
    """SynthesisLayer.forward, networks.py:286-305."""
    cout, cin, k, _ = sd[p + "weight"].shape
    assert tuple(x.shape[1:]) == (cin, res // up, res // up)
    styles = fully_connected(w, sd[p + "affine.weight"], sd[p + "affine.bias"])
    noise = _noise_for(sd, p, x.shape[0], res, noise_mode, noises, x.device) if use_noise else None
    x = R.modulated_conv2d(x, sd[p + "weight"], styles, noise=noise, up=up, padding=k // 2,
                           resample_filter=sd[p + "resample_filter"], flip_weight=(up == 1),
                           fused_modconv=fused_modconv)
    act_gain = R.act_defaults(activation)[1] * gain
    act_clamp = conv_clamp * gain if conv_clamp is not None else None
    return R.bias_act(x, sd[p + "bias"].to(x.dtype), act=activation, gain=act_gain, clamp=act_clamp)


def torgb_layer(x, w, sd, p, fused_modconv, conv_clamp=None):
    """ToRGBLayer.forward, networks.py:320-324: 1x1, no demodulation, linear bias_act."""
    cout, cin, k, _ = sd[p + "weight"].shape
    styles = fully_connected(w, sd[p + "affine.weight"], sd[p + "affine.bias"]) * (1 / math.sqrt(cin * k * k))
    x = R.modulated_conv2d(x, sd[p + "weight"], styles, demodulate=False, fused_modconv=fused_modconv)
    return R.bias_act(x, sd[p + "bias"].to(x.dtype), clamp=conv_clamp)


def conv2d_layer(x, sd, p, up=1, down=1, activation="linear", gain=1.0, conv_clamp=None):
    """Conv2dLayer.forward, networks.py:158-169 (only used by the 'resnet' skip branch)."""
    wt = sd[p + "weight"]
    cout, cin, k, _ = wt.shape
    w = wt * (1 / math.sqrt(cin * k * k))
    b = sd.get(p + "bias")
    x = R.conv2d_resample(x, w.to(x.dtype), f=sd[p + "resample_filter"], up=up, down=down, padding=k // 2,
                          flip_weight=(up == 1))
    act_gain = R.act_defaults(activation)[1] * gain
    act_clamp = conv_clamp * gain if conv_clamp is not None else None
    return R.bias_act(x, b.to(x.dtype) if b is not None else None, act=activation, gain=act_gain, clamp=act_clamp)


def synthesis_block(x, img, ws, sd, p, res, first, is_last, training, noise_mode="random", fused_modconv=None,
                    architecture="skip", conv_clamp=None, noises=None, dtype=torch.float32):
    """SynthesisBlock.forward, networks.py:379-419 (fp32 / contiguous layout)."""
    w_iter = iter(ws.unbind(dim=1))
    if fused_modconv is None:  # :384-386
        fused_modconv = (not training) and (dtype == torch.float32 or int(ws.shape[0]) == 1)
    kw = dict(noise_mode=noise_mode, fused_modconv=fused_modconv, conv_clamp=conv_clamp, noises=noises)

    if first:  # :389-391
        x = sd[p + "const"].to(dtype).unsqueeze(0).repeat([ws.shape[0], 1, 1, 1])
        x = synthesis_layer(x, next(w_iter), sd, p + "conv1.", res, 1, **kw)
    elif architecture == "resnet":  # :399-403
        y = conv2d_layer(x.to(dtype), sd, p + "skip.", up=2, gain=math.sqrt(0.5))
        x = synthesis_layer(x.to(dtype), next(w_iter), sd, p + "conv0.", res, 2, **kw)
        x = synthesis_layer(x, next(w_iter), sd, p + "conv1.", res, 1, gain=math.sqrt(0.5), **kw)
        x = y + x
    else:  # :405-406
        x = synthesis_layer(x.to(dtype), next(w_iter), sd, p + "conv0.", res, 2, **kw)
        x = synthesis_layer(x, next(w_iter), sd, p + "conv1.", res, 1, **kw)

    if img is not None:  # :409-411
        img = R.upsample2d(img, sd[p + "resample_filter"])
    if is_last or architecture == "skip":  # :412-415
        y = torgb_layer(x, next(w_iter), sd, p + "torgb.", fused_modconv, conv_clamp=conv_clamp)
        y = y.to(torch.float32)
        img = img + y if img is not None else y
    return x, img


def synthesis_network(ws, sd, img_resolution, prefix="", training=False, **block_kwargs):
    """SynthesisNetwork.forward, networks.py:457-472 (ws split with torgb overlap :463-466)."""
    log2 = int(math.log2(img_resolution))
    resolutions = [2 ** i for i in range(2, log2 + 1)]
    architecture = block_kwargs.get("architecture", "skip")
    assert tuple(ws.shape[1:]) == (synthesis_num_ws(img_resolution), ws.shape[2])
    ws = ws.to(torch.float32)
    x = img = None
    w_idx = 0
    for res in resolutions:
        first, is_last = res == 4, res == img_resolution
        num_conv = 1 if first else 2
        num_torgb = 1 if (is_last or architecture == "skip") else 0
        cur = ws.narrow(1, w_idx, num_conv + num_torgb)
        w_idx += num_conv
        x, img = synthesis_block(x, img, cur, sd, f"{prefix}b{res}.", res, first, is_last, training, **block_kwargs)
    return img


# ---------------------------------------------------------------------------
# mapping side
# ---------------------------------------------------------------------------


def mapping_network(z, sd, prefix, num_layers, num_ws, lr_multiplier=0.01, truncation_psi=1.0,
                    truncation_cutoff=None):
    """MappingNetwork.forward with c_dim = 0, networks.py:213-249 (no w_avg update: eval mode,
    which is how VisCoIN runs the original generator - SURVEY D.3)."""
    x = normalize_2nd_moment(z.to(torch.float32))
    for i in range(num_layers):
        x = fully_connected(x, sd[f"{prefix}fc{i}.weight"], sd[f"{prefix}fc{i}.bias"], activation="lrelu",
                            lr_multiplier=lr_multiplier)
    if num_ws is not None:
        x = x.unsqueeze(1).repeat([1, num_ws, 1])
    if truncation_psi != 1:
        w_avg = sd[prefix + "w_avg"]
        if num_ws is None or truncation_cutoff is None:
            x = w_avg.lerp(x, truncation_psi)
        else:
            x = torch.cat([w_avg.lerp(x[:, :truncation_cutoff], truncation_psi), x[:, truncation_cutoff:]], dim=1)
    return x


def generator(z, sd, img_resolution, mapping_layers=8, training=False, truncation_psi=1.0, **synthesis_kwargs):
    """Generator.forward, networks.py:497-500."""
    ws = mapping_network(z, sd, "mapping.", mapping_layers, synthesis_num_ws(img_resolution),
                         truncation_psi=truncation_psi)
    return synthesis_network(ws, sd, img_resolution, prefix="synthesis.", training=training, **synthesis_kwargs)


def mapping_adapted(z1, z2, sd, prefix, num_ws, z1_dim, z2_dim, num_layers=1, coarse_layer=2, mid_layer=10,
                    lr_multiplier=0.01):
    """MappingNetworkAdapted.forward, viscoin/models/gan.py:211-234: one FC stack per w;
    stacks coarse..mid read z1, the others z2; fixed_w_avg added to every w."""
    if z1.ndim > 3:
        z1 = z1.reshape(-1, z1_dim)
    assert z1.shape[1] == z1_dim and z2.shape[1] == z2_dim
    x1 = normalize_2nd_moment(z1.to(torch.float32))
    x2 = normalize_2nd_moment(z2.to(torch.float32))
    outs = []
    for i in range(num_ws):
        x = x2 if (i < coarse_layer or i > mid_layer) else x1
        for j in range(num_layers):
            x = fully_connected(x, sd[f"{prefix}s{i}.{j}.weight"], sd[f"{prefix}s{i}.{j}.bias"],
                                activation="lrelu", lr_multiplier=lr_multiplier)
        outs.append(x)
    ws = torch.stack(outs, dim=1)
    return ws + sd[prefix + "fixed_w_avg"].reshape(1, 1, -1)


def generator_adapted(z1, z2, sd, img_resolution=256, z_dim=256, training=False, return_latents=False,
                      **synthesis_kwargs):
    """GeneratorAdapted.forward, viscoin/models/gan.py:85-105 (low_res256 False)."""
    ws = mapping_adapted(z1, z2, sd, "mapping.", synthesis_num_ws(img_resolution), 9 * z_dim, 9 * z_dim)
    img = synthesis_network(ws, sd, img_resolution, prefix="synthesis.", training=training, **synthesis_kwargs)
    return (img, ws) if return_latents else img
