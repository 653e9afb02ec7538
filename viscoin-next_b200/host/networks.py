"""networks - the consumers of the hot path, mirrored from the reference with identical parameter / buffer names so
that reference ``state_dict``s load unchanged:

  FullyConnectedLayer   stylegan2_ada/training/networks.py:89-118
  SynthesisLayer        :254-305     ToRGBLayer :310-324     SynthesisBlock :329-419     SynthesisNetwork :424-472
  MappingNetwork        :174-249 (c_dim = 0)                 Generator :477-500
  MappingNetworkAdapted / GeneratorAdapted                   viscoin/models/gan.py:35-234

Every layer body is ONE fused call into the sm_100a library (``modconv.modconv_layer``: modulation -> tcgen05
implicit GEMM -> demodulation + noise + bias + activation + clamp in the epilogue); the RGB skip uses
``upfirdn2d.upsample2d``; mapping FCs use ``bias_act``.  The small dense FCs (affine 512->C, mapping) are plain
library GEMMs (torch.addmm / matmul).  All blocks run in fp32 storage (``force_fp32`` semantics of :382-383); the
convolution operands are bf16 on the tensor cores (tolerances in tests/).
"""

import math
import os

import numpy as np
import torch

from . import bias_act as _bias_act
from .misc import profiled_function as _profiled
from . import modconv as _modconv
from . import upfirdn2d as _upfirdn2d


@_profiled
def normalize_2nd_moment(x, dim=1, eps=1e-8):
    return x * (x.square().mean(dim=dim, keepdim=True) + eps).rsqrt()


rgb_tap = os.environ.get("VCN_RGB_TAP", "1") == "1"


class FullyConnectedLayer(torch.nn.Module):
    def __init__(self, in_features, out_features, bias=True, activation="linear", lr_multiplier=1, bias_init=0):
        super().__init__()
        self.activation = activation
        self.weight = torch.nn.Parameter(torch.randn([out_features, in_features]) / lr_multiplier)
        self.bias = torch.nn.Parameter(torch.full([out_features], np.float32(bias_init))) if bias else None
        self.weight_gain = lr_multiplier / math.sqrt(in_features)
        self.bias_gain = lr_multiplier

    def forward(self, x):
        w = self.weight.to(x.dtype) * self.weight_gain
        b = self.bias
        if b is not None:
            b = b.to(x.dtype)
            if self.bias_gain != 1:
                b = b * self.bias_gain
        if self.activation == "linear" and b is not None:
            return torch.addmm(b.unsqueeze(0), x, w.t())
        x = x.matmul(w.t())
        return _bias_act.bias_act(x, b, act=self.activation)


class SynthesisLayer(torch.nn.Module):
    def __init__(self, in_channels, out_channels, w_dim, resolution, kernel_size=3, up=1, use_noise=True,
                 activation="lrelu", resample_filter=[1, 3, 3, 1], conv_clamp=None, channels_last=False):
        super().__init__()
        self.resolution, self.up, self.use_noise, self.activation, self.conv_clamp = resolution, up, use_noise, activation, conv_clamp
        self.register_buffer("resample_filter", _upfirdn2d.setup_filter(resample_filter))
        self.padding = kernel_size // 2
        self.act_gain = _bias_act.activation_funcs[activation].def_gain
        self.affine = FullyConnectedLayer(w_dim, in_channels, bias_init=1)
        self.weight = torch.nn.Parameter(torch.randn([out_channels, in_channels, kernel_size, kernel_size]))
        if use_noise:
            self.register_buffer("noise_const", torch.randn([resolution, resolution]))
            self.noise_strength = torch.nn.Parameter(torch.zeros([]))
        self.bias = torch.nn.Parameter(torch.zeros([out_channels]))

    def _const_noise(self):
        """noise_const * noise_strength (networks.py:296); cached per (buffer, parameter) version while the strength is frozen."""
        ns = self.noise_strength
        if ns.requires_grad and torch.is_grad_enabled():
            return self.noise_const * ns
        key = (self.noise_const.data_ptr(), self.noise_const._version, ns.data_ptr(), ns._version)
        hit = getattr(self, "_noise_cache", None)
        if hit is None or hit[0] != key:
            if ns.is_cuda and torch.cuda.is_current_stream_capturing():
                return self.noise_const * ns  # a tensor born inside a capture belongs to that graph's pool: do not keep it
            with torch.no_grad():
                hit = (key, self.noise_const * ns)
            self._noise_cache = hit
        return hit[1]

    def forward(self, x, w, noise_mode="random", fused_modconv=True, gain=1, styles=None):
        assert noise_mode in ["random", "const", "none"]
        in_res = self.resolution // self.up
        assert tuple(x.shape[1:]) == (self.weight.shape[1], in_res, in_res)
        if styles is None:  # else: precomputed by SynthesisNetwork._batched_styles (all affine FCs in one batched GEMM)
            styles = self.affine(w)
        noise = None
        if self.use_noise and noise_mode == "random":
            noise = torch.randn([x.shape[0], 1, self.resolution, self.resolution], device=x.device) * self.noise_strength
        if self.use_noise and noise_mode == "const":
            noise = self._const_noise()
        act_clamp = self.conv_clamp * gain if self.conv_clamp is not None else None
        return _modconv.modconv_layer(x, self.weight, styles, noise=noise, bias=self.bias, up=self.up, padding=self.padding,
                                      resample_filter=self.resample_filter, flip_weight=(self.up == 1),
                                      act=self.activation, gain=self.act_gain * gain, clamp=act_clamp)

    def forward_nhwc(self, act, w, noise_mode="random", gain=1, styles=None):
        """Resident-activation (bf16 NHWC) variant of forward(); inference only."""
        if styles is None:
            styles = self.affine(w)
        noise = None
        if self.use_noise and noise_mode == "random":
            noise = torch.randn([act.shape[0], 1, self.resolution, self.resolution], device=act.device) * self.noise_strength
        if self.use_noise and noise_mode == "const":
            noise = self._const_noise()
        act_clamp = self.conv_clamp * gain if self.conv_clamp is not None else None
        return _modconv.layer_nhwc(act, self.weight, styles, noise=noise, bias=self.bias, up=self.up,
                                   resample_filter=self.resample_filter, act_name=self.activation,
                                   gain=self.act_gain * gain, clamp=act_clamp)


class ToRGBLayer(torch.nn.Module):
    def __init__(self, in_channels, out_channels, w_dim, kernel_size=1, conv_clamp=None, channels_last=False):
        super().__init__()
        self.conv_clamp = conv_clamp
        self.affine = FullyConnectedLayer(w_dim, in_channels, bias_init=1)
        self.weight = torch.nn.Parameter(torch.randn([out_channels, in_channels, kernel_size, kernel_size]))
        self.bias = torch.nn.Parameter(torch.zeros([out_channels]))
        self.weight_gain = 1 / math.sqrt(in_channels * (kernel_size ** 2))

    def forward(self, x, w, fused_modconv=True, styles=None, tap=False):
        """tap=True: returns (rgb, x2) with x2 a second handle on x for the next block (modconv.modconv_layer)."""
        if styles is None:
            styles = self.affine(w) * self.weight_gain
        return _modconv.modconv_layer(x, self.weight, styles, bias=self.bias, demodulate=False, act="linear",
                                      clamp=self.conv_clamp, tap=tap)

    def forward_nhwc(self, act, w, styles=None):
        """bf16 NHWC activation -> fp32 NCHW RGB: per-sample 1x1 weights, no pass over the activation besides the GEMM."""
        if styles is None:
            styles = self.affine(w) * self.weight_gain
        return _modconv.layer_nhwc(act, self.weight, styles, bias=self.bias, demodulate=False, act_name="linear",
                                   clamp=self.conv_clamp, per_sample=True, out_nchw_f32=True)


class Conv2dLayer(torch.nn.Module):
    """networks.py:123-169, the configurations the synthesis side uses (the 'resnet' skip: 1x1, up=2, no bias)."""

    def __init__(self, in_channels, out_channels, kernel_size, bias=True, activation="linear", up=1, down=1,
                 resample_filter=[1, 3, 3, 1], conv_clamp=None, channels_last=False, trainable=True):
        super().__init__()
        assert down == 1 and kernel_size in (1, 3), "only the synthesis-side configurations are implemented"
        self.activation, self.up, self.conv_clamp = activation, up, conv_clamp
        self.register_buffer("resample_filter", _upfirdn2d.setup_filter(resample_filter))
        self.padding = kernel_size // 2
        self.weight_gain = 1 / math.sqrt(in_channels * (kernel_size ** 2))
        self.act_gain = _bias_act.activation_funcs[activation].def_gain
        self.weight = torch.nn.Parameter(torch.randn([out_channels, in_channels, kernel_size, kernel_size]))
        self.bias = torch.nn.Parameter(torch.zeros([out_channels])) if bias else None

    def forward(self, x, gain=1):
        w = self.weight * self.weight_gain
        act_clamp = self.conv_clamp * gain if self.conv_clamp is not None else None
        if self.up == 1 or self.weight.shape[-1] == 3:
            return _modconv.modconv_layer(x, w, None, bias=self.bias, up=self.up, resample_filter=self.resample_filter,
                                          demodulate=False, flip_weight=(self.up == 1), act=self.activation,
                                          gain=self.act_gain * gain, clamp=act_clamp)
        # 1x1 with up=2: convolve at low resolution, then FIR-upsample (conv2d_resample.py:107-110)
        x = _modconv.modconv_layer(x, w, None, demodulate=False, act=None)
        x = _upfirdn2d.upsample2d(x, self.resample_filter)
        return _bias_act.bias_act(x, self.bias, act=self.activation, gain=self.act_gain * gain, clamp=act_clamp)


class SynthesisBlock(torch.nn.Module):
    def __init__(self, in_channels, out_channels, w_dim, resolution, img_channels, is_last, architecture="skip",
                 resample_filter=[1, 3, 3, 1], conv_clamp=None, use_fp16=False, fp16_channels_last=False, **layer_kwargs):
        assert architecture in ["orig", "skip", "resnet"]
        super().__init__()
        self.in_channels, self.w_dim, self.resolution, self.img_channels = in_channels, w_dim, resolution, img_channels
        self.is_last, self.architecture = is_last, architecture
        # networks.py:382-383,394: a use_fp16 block stores its activations in 16 bits, channels_last.  Here that is the bf16
        # resident pipeline (bf16 has fp32's exponent range: the fp16 pre-normalisation of networks.py:47-49 is not needed;
        # conv_clamp is applied in the epilogue exactly as in the reference)
        self.use_fp16 = bool(use_fp16)
        self.register_buffer("resample_filter", _upfirdn2d.setup_filter(resample_filter))
        self.num_conv = self.num_torgb = 0
        if in_channels == 0:
            self.const = torch.nn.Parameter(torch.randn([out_channels, resolution, resolution]))
        if in_channels != 0:
            self.conv0 = SynthesisLayer(in_channels, out_channels, w_dim=w_dim, resolution=resolution, up=2,
                                        resample_filter=resample_filter, conv_clamp=conv_clamp, **layer_kwargs)
            self.num_conv += 1
        self.conv1 = SynthesisLayer(out_channels, out_channels, w_dim=w_dim, resolution=resolution, conv_clamp=conv_clamp,
                                    **layer_kwargs)
        self.num_conv += 1
        if is_last or architecture == "skip":
            self.torgb = ToRGBLayer(out_channels, img_channels, w_dim=w_dim, conv_clamp=conv_clamp)
            self.num_torgb += 1
        if in_channels != 0 and architecture == "resnet":
            self.skip = Conv2dLayer(in_channels, out_channels, kernel_size=1, bias=False, up=2, resample_filter=resample_filter)

    def forward_nhwc(self, act, img, ws, noise_mode="random", styles=None):
        """Resident-activation block ('skip' / 'orig' architectures): act stays bf16 NHWC, img fp32 NCHW."""
        w_iter = iter(ws.unbind(dim=1))
        st = styles or {}
        if self.in_channels == 0:
            act = self.const.to(torch.bfloat16).permute(1, 2, 0).unsqueeze(0).repeat([ws.shape[0], 1, 1, 1]).contiguous()
            act = self.conv1.forward_nhwc(act, next(w_iter), noise_mode=noise_mode, styles=st.get(self.conv1))
        else:
            act = self.conv0.forward_nhwc(act, next(w_iter), noise_mode=noise_mode, styles=st.get(self.conv0))
            act = self.conv1.forward_nhwc(act, next(w_iter), noise_mode=noise_mode, styles=st.get(self.conv1))
        if img is not None:
            img = _upfirdn2d.upsample2d(img, self.resample_filter)
        if self.is_last or self.architecture == "skip":
            y = self.torgb.forward_nhwc(act, next(w_iter), styles=st.get(self.torgb))
            img = img.add_(y) if img is not None else y
        return act, img

    def forward(self, x, img, ws, force_fp32=False, fused_modconv=None, resident=False, styles=None, **layer_kwargs):
        """resident=True (training on the 'skip' / 'orig' architectures): activations are bf16 channels_last tensors from
        block to block, forward and backward (modconv._ModConvLayer keeps them that way); img stays fp32 NCHW."""
        assert tuple(ws.shape[1:]) == (self.num_conv + self.num_torgb, self.w_dim)
        w_iter = iter(ws.unbind(dim=1))
        resident = (resident or (self.use_fp16 and not force_fp32)) and self.architecture != "resnet"
        if self.in_channels == 0:
            x = self.const.to(torch.float32).unsqueeze(0).repeat([ws.shape[0], 1, 1, 1])
            if resident:
                x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        elif not resident:
            x = x.to(torch.float32)
        elif x.dtype != torch.bfloat16 or x.stride(1) != 1:  # first 16-bit block after fp32 ones (networks.py:394)
            x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        st = styles or {}
        if self.in_channels == 0:
            x = self.conv1(x, next(w_iter), styles=st.get(self.conv1), **layer_kwargs)
        elif self.architecture == "resnet":
            y = self.skip(x, gain=math.sqrt(0.5))
            x = self.conv0(x, next(w_iter), styles=st.get(self.conv0), **layer_kwargs)
            x = self.conv1(x, next(w_iter), gain=math.sqrt(0.5), styles=st.get(self.conv1), **layer_kwargs)
            x = y + x
        else:
            x = self.conv0(x, next(w_iter), styles=st.get(self.conv0), **layer_kwargs)
            x = self.conv1(x, next(w_iter), styles=st.get(self.conv1), **layer_kwargs)
        if img is not None:
            img = _upfirdn2d.upsample2d(img, self.resample_filter)
        if self.is_last or self.architecture == "skip":
            # training on the resident pipeline: x is read by ToRGB and by the next block - the ToRGB layer hands back a second
            # handle on x so that both gradients meet inside its fused backward (one pass instead of two + a torch add)
            tap = (rgb_tap and resident and not self.is_last and torch.is_grad_enabled() and x.requires_grad
                   and x.dtype == torch.bfloat16 and _modconv.torgb_fused)
            if tap:
                y, x = self.torgb(x, next(w_iter), styles=st.get(self.torgb), tap=True)
                y = y.to(torch.float32)
            else:
                y = self.torgb(x, next(w_iter), styles=st.get(self.torgb)).to(torch.float32)
            img = img.add_(y) if img is not None else y
        return x, img


class SynthesisNetwork(torch.nn.Module):
    def __init__(self, w_dim, img_resolution, img_channels, channel_base=32768, channel_max=512, num_fp16_res=0,
                 **block_kwargs):
        assert img_resolution >= 4 and img_resolution & (img_resolution - 1) == 0
        super().__init__()
        self.w_dim, self.img_resolution, self.img_channels = w_dim, img_resolution, img_channels
        self.img_resolution_log2 = int(np.log2(img_resolution))
        self.num_fp16_res = num_fp16_res
        # differentiable forward with bf16 channels_last activations between layers (see SynthesisBlock.forward); off by
        # default: the reference's configurations store fp32 (networks.py:382-383 with num_fp16_res = 0)
        self.resident_bf16 = False
        self.batch_affine = True  # all style FCs of a forward in one batched GEMM while their parameters are frozen
        self.block_resolutions = [2 ** i for i in range(2, self.img_resolution_log2 + 1)]
        channels_dict = {res: min(channel_base // res, channel_max) for res in self.block_resolutions}
        fp16_resolution = max(2 ** (self.img_resolution_log2 + 1 - num_fp16_res), 8)  # networks.py:442
        self.num_ws = 0
        for res in self.block_resolutions:
            in_channels = channels_dict[res // 2] if res > 4 else 0
            block = SynthesisBlock(in_channels, channels_dict[res], w_dim=w_dim, resolution=res, img_channels=img_channels,
                                   is_last=(res == img_resolution), use_fp16=(res >= fp16_resolution), **block_kwargs)
            self.num_ws += block.num_conv
            if res == img_resolution:
                self.num_ws += block.num_torgb
            setattr(self, f"b{res}", block)

    def _affine_plan(self):
        """(layer, index into ws, extra gain) of every style FC in forward order (networks.py:463-466 ws split; ToRGB
        multiplies its styles by weight_gain, networks.py:321)."""
        plan, w_idx = [], 0
        for res in self.block_resolutions:
            b = getattr(self, f"b{res}")
            k = w_idx
            if b.in_channels != 0:
                plan.append((b.conv0, k, 1.0)); k += 1
            plan.append((b.conv1, k, 1.0)); k += 1
            if b.is_last or b.architecture == "skip":
                plan.append((b.torgb, k, b.torgb.weight_gain))
            w_idx += b.num_conv
        return plan

    def _batched_styles(self, ws):
        """All style FCs of one forward (20 at 256^2: `affine` of every SynthesisLayer / ToRGBLayer, networks.py:290,321) as ONE
        batched GEMM  S[l] = ws[:, idx_l] @ (W_l g_l)^T + b_l g_l  over zero-padded [L, w_dim, Cmax] weights (SURVEY 8(f) N1:
        ~45 tiny launches per forward become 3).  Only while the affine parameters are frozen (VisCoIN's shared synthesis
        network, inference): the padded stack is cached per parameter version; ws may require grad (baddbmm's autograd).
        Returns {layer: styles [N, Cin_l]} or None when the per-layer path must run."""
        plan = self._affine_plan()
        params = [p for layer, _, _ in plan for p in (layer.affine.weight, layer.affine.bias)]
        if not ws.is_cuda or (torch.is_grad_enabled() and any(p.requires_grad for p in params)):
            return None
        key = tuple((p.data_ptr(), p._version) for p in params) + (ws.device,)
        cache = getattr(self, "_affine_cache", None)
        if cache is None or cache[0] != key:
            if torch.cuda.is_current_stream_capturing():
                return None  # build the stack eagerly (warm-up), never inside a capture
            with torch.no_grad():
                cmax = max(layer.affine.weight.shape[0] for layer, _, _ in plan)
                wt = torch.zeros([len(plan), self.w_dim, cmax], dtype=torch.float32, device=ws.device)
                bs = torch.zeros([len(plan), 1, cmax], dtype=torch.float32, device=ws.device)
                for l, (layer, _, g) in enumerate(plan):
                    a = layer.affine
                    c = a.weight.shape[0]
                    wt[l, :, :c] = (a.weight.float() * (a.weight_gain * g)).t()
                    bs[l, 0, :c] = a.bias.float() * (a.bias_gain * g)
                idx = torch.tensor([k for _, k, _ in plan], dtype=torch.long, device=ws.device)
            cache = (key, wt, bs, idx)
            self._affine_cache = cache
        _, wt, bs, idx = cache
        s_all = torch.baddbmm(bs, ws.index_select(1, idx).transpose(0, 1), wt)  # [L, N, Cmax]
        return {layer: s_all[l, :, :layer.affine.weight.shape[0]] for l, (layer, _, _) in enumerate(plan)}

    def forward(self, ws, **block_kwargs):
        assert tuple(ws.shape[1:]) == (self.num_ws, self.w_dim)
        ws = ws.to(torch.float32)
        styles = self._batched_styles(ws) if self.batch_affine else None
        block_ws, w_idx = [], 0
        for res in self.block_resolutions:
            block = getattr(self, f"b{res}")
            block_ws.append(ws.narrow(1, w_idx, block.num_conv + block.num_torgb))
            w_idx += block.num_conv
        x = img = None
        # resident_bf16 and nothing to differentiate (no_grad, or frozen weights and a constant ws): the bf16-NHWC inference
        # pipeline.  With resident_bf16 = False the same model gives the same (fp32-storage) numerics with and without grad.
        no_grad = (not torch.is_grad_enabled()) or not (ws.requires_grad or any(p.requires_grad for p in self.parameters()))
        fast = self.resident_bf16 and no_grad and ws.is_cuda and set(block_kwargs) <= {"noise_mode"} \
            and all(getattr(self, f"b{res}").architecture in ("skip", "orig") for res in self.block_resolutions)
        for res, cur_ws in zip(self.block_resolutions, block_ws):
            if fast:  # inference: activations stay resident as bf16 NHWC, no per-layer layout / modulation passes
                x, img = getattr(self, f"b{res}").forward_nhwc(x, img, cur_ws, styles=styles, **block_kwargs)
            else:
                x, img = getattr(self, f"b{res}")(x, img, cur_ws, resident=self.resident_bf16, styles=styles, **block_kwargs)
        return img


class MappingNetwork(torch.nn.Module):
    def __init__(self, z_dim, c_dim, w_dim, num_ws, num_layers=8, embed_features=None, layer_features=None,
                 activation="lrelu", lr_multiplier=0.01, w_avg_beta=0.995):
        super().__init__()
        assert c_dim == 0, "conditional mapping is not on the VisCoIN path"
        self.z_dim, self.c_dim, self.w_dim, self.num_ws, self.num_layers, self.w_avg_beta = z_dim, c_dim, w_dim, num_ws, num_layers, w_avg_beta
        if layer_features is None:
            layer_features = w_dim
        features = [z_dim] + [layer_features] * (num_layers - 1) + [w_dim]
        for idx in range(num_layers):
            setattr(self, f"fc{idx}", FullyConnectedLayer(features[idx], features[idx + 1], activation=activation,
                                                         lr_multiplier=lr_multiplier))
        if num_ws is not None and w_avg_beta is not None:
            self.register_buffer("w_avg", torch.zeros([w_dim]))

    def forward(self, z, c=None, truncation_psi=1, truncation_cutoff=None, skip_w_avg_update=False):
        x = normalize_2nd_moment(z.to(torch.float32))
        for idx in range(self.num_layers):
            x = getattr(self, f"fc{idx}")(x)
        if self.w_avg_beta is not None and self.training and not skip_w_avg_update:
            self.w_avg.copy_(x.detach().mean(dim=0).lerp(self.w_avg, self.w_avg_beta))
        if self.num_ws is not None:
            x = x.unsqueeze(1).repeat([1, self.num_ws, 1])
        if truncation_psi != 1:
            if self.num_ws is None or truncation_cutoff is None:
                x = self.w_avg.lerp(x, truncation_psi)
            else:
                x[:, :truncation_cutoff] = self.w_avg.lerp(x[:, :truncation_cutoff], truncation_psi)
        return x


class Generator(torch.nn.Module):
    def __init__(self, z_dim, c_dim, w_dim, img_resolution, img_channels, mapping_kwargs={}, synthesis_kwargs={}):
        super().__init__()
        self.z_dim, self.c_dim, self.w_dim, self.img_resolution, self.img_channels = z_dim, c_dim, w_dim, img_resolution, img_channels
        self.synthesis = SynthesisNetwork(w_dim=w_dim, img_resolution=img_resolution, img_channels=img_channels, **synthesis_kwargs)
        self.num_ws = self.synthesis.num_ws
        self.mapping = MappingNetwork(z_dim=z_dim, c_dim=c_dim, w_dim=w_dim, num_ws=self.num_ws, **mapping_kwargs)

    def forward(self, z, c=None, truncation_psi=1, truncation_cutoff=None, **synthesis_kwargs):
        ws = self.mapping(z, c, truncation_psi=truncation_psi, truncation_cutoff=truncation_cutoff)
        return self.synthesis(ws, **synthesis_kwargs)


class _StackedParams(torch.autograd.Function):
    """Identity on `stacked` whose gradient is routed, slice by slice, to the parameters that alias its storage."""

    @staticmethod
    def forward(ctx, stacked, *params):
        ctx.n = len(params)
        return stacked.view_as(stacked)

    @staticmethod
    def backward(ctx, g):
        return (None,) + tuple(g[i] for i in range(ctx.n))


class MappingNetworkAdapted(torch.nn.Module):
    """viscoin/models/gan.py:147-234: one FC stack per w; stacks coarse..mid read z1, the others z2.

    SURVEY 8(f) N1: with one layer per stack (VisCoIN's default) the 14 FCs (2304 -> 512, lrelu) run as TWO batched GEMMs
    (the z2-fed stacks 0,1,11,12,13 and the z1-fed stacks 2..10) and ONE ``bias_act`` launch over [14*N, 512] instead of
    14 GEMMs + 14 ``bias_act`` launches - forward, input gradient and weight gradient alike (baddbmm's autograd is two
    batched GEMMs).  The 14 weight Parameters keep their names (reference ``state_dict``s load unchanged) but alias one
    stacked [14, 512, 2304] buffer, so no per-step gather copy exists: Adam updates write through to it."""

    def __init__(self, z1_dim, z2_dim, w_dim, num_ws, num_layers=1, coarse_layer=2, mid_layer=10, activation="lrelu",
                 lr_multiplier=0.01):
        super().__init__()
        self.z1_dim, self.z2_dim, self.w_dim, self.num_ws, self.num_layers = z1_dim, z2_dim, w_dim, num_ws, num_layers
        self.coarse_layer, self.mid_layer = coarse_layer, mid_layer
        self.fixed_w_avg = torch.nn.Parameter(torch.zeros([w_dim]))
        for i in range(num_ws):
            z_dim = z2_dim if (i < coarse_layer or i > mid_layer) else z1_dim
            feats = [z_dim] + [w_dim] * num_layers
            stack = torch.nn.ModuleList([FullyConnectedLayer(feats[j], feats[j + 1], activation=activation,
                                                             lr_multiplier=lr_multiplier) for j in range(num_layers)])
            setattr(self, f"s{i}", stack)

    grouped = True  # batched-GEMM path on CUDA tensors (False: the per-stack loop of the reference)

    def _stacked(self):
        """(W [L,512,K], B [L,512], order): the stacked buffers the layer Parameters alias, z2-fed stacks first.  Re-built
        (and the Parameters re-pointed) whenever a Parameter's storage moved (``.to()``, a replaced ``.data``)."""
        order = [i for i in range(self.num_ws) if (i < self.coarse_layer or i > self.mid_layer)]
        n2 = len(order)
        order += [i for i in range(self.num_ws) if not (i < self.coarse_layer or i > self.mid_layer)]
        layers = [getattr(self, f"s{i}")[0] for i in order]
        hit = getattr(self, "_stack_cache", None)
        ok = hit is not None and all(l.weight.data_ptr() == hit[0][j].data_ptr() and l.bias.data_ptr() == hit[1][j].data_ptr()
                                     and l.weight.device == hit[0].device for j, l in enumerate(layers))
        if not ok:
            with torch.no_grad():
                W = torch.stack([l.weight.detach() for l in layers])
                B = torch.stack([l.bias.detach() for l in layers])
                for j, l in enumerate(layers):
                    l.weight.data = W[j]
                    l.bias.data = B[j]
            hit = (W, B, order, n2)
            self._stack_cache = hit
        return hit + (layers,)

    def _forward_grouped(self, x1, x2):
        W, B, order, n2, layers = self._stacked()
        l0 = layers[0]
        Wg = _StackedParams.apply(W, *[l.weight for l in layers])
        Bg = _StackedParams.apply(B, *[l.bias for l in layers])
        n = x1.shape[0]
        bias = Bg.unsqueeze(1) * l0.bias_gain if l0.bias_gain != 1 else Bg.unsqueeze(1)
        y2 = torch.baddbmm(bias[:n2], x2.unsqueeze(0).expand(n2, n, x2.shape[1]), Wg[:n2].transpose(1, 2), alpha=l0.weight_gain)
        y1 = torch.baddbmm(bias[n2:], x1.unsqueeze(0).expand(len(order) - n2, n, x1.shape[1]), Wg[n2:].transpose(1, 2), alpha=l0.weight_gain)
        y = _bias_act.bias_act(torch.cat([y2, y1]).reshape(len(order) * n, self.w_dim), None, act=l0.activation)
        key = (tuple(order), y.device)
        if getattr(self, "_inv_cache", (None,))[0] != key:
            inv = torch.empty(len(order), dtype=torch.long)
            inv[torch.tensor(order)] = torch.arange(len(order))  # y row j holds stack order[j]
            self._inv_cache = (key, inv.to(y.device))
        return y.reshape(len(order), n, self.w_dim).index_select(0, self._inv_cache[1]).transpose(0, 1) + self.fixed_w_avg

    def forward(self, z1, z2):
        if z1.ndim > 3:
            z1 = z1.reshape(-1, self.z1_dim)
        x1 = normalize_2nd_moment(z1.to(torch.float32))
        x2 = normalize_2nd_moment(z2.to(torch.float32))
        if self.grouped and x1.is_cuda and self.num_layers == 1 and self.z1_dim == self.z2_dim and self.s0[0].bias is not None:
            return self._forward_grouped(x1, x2)
        outs = []
        for i in range(self.num_ws):
            x = x2 if (i < self.coarse_layer or i > self.mid_layer) else x1
            for layer in getattr(self, f"s{i}"):
                x = layer(x)
            outs.append(x)
        return torch.stack(outs, dim=1) + self.fixed_w_avg


class GeneratorAdapted(torch.nn.Module):
    """viscoin/models/gan.py:35-144 (low_res256 False)."""

    def __init__(self, z_dim=256, w_dim=512, img_resolution=256, img_channels=3, mapping_kwargs={}, synthesis_kwargs={}):
        super().__init__()
        self.z_dim, self.w_dim, self.img_resolution, self.img_channels = z_dim, w_dim, img_resolution, img_channels
        self.synthesis = SynthesisNetwork(w_dim=w_dim, img_resolution=img_resolution, img_channels=img_channels, **synthesis_kwargs)
        self.num_ws = self.synthesis.num_ws
        self.mapping = MappingNetworkAdapted(9 * z_dim, 9 * z_dim, w_dim, self.num_ws, **mapping_kwargs)

    @classmethod
    def from_gan(cls, gan, z_dim=256, w_dim=512, mapping_kwargs={}):
        assert gan.synthesis is not None, "The provided GAN is not a valid StyleGAN model"
        model = cls(z_dim=z_dim, w_dim=w_dim, mapping_kwargs=mapping_kwargs)
        model.synthesis = gan.synthesis  # the SAME module object is shared by both generators (gan.py:142)
        model.num_ws, model.img_resolution, model.img_channels = gan.synthesis.num_ws, gan.synthesis.img_resolution, gan.synthesis.img_channels
        return model

    def forward(self, z1, z2, return_latents=False, truncation_psi=1, truncation_cutoff=None, **synthesis_kwargs):
        ws = self.mapping(z1, z2)
        img = self.synthesis(ws, **synthesis_kwargs)
        return (img, ws) if return_latents else img

    def gen_from_w(self, w, **synthesis_kwargs):
        return self.synthesis(w, **synthesis_kwargs)


class GraphedSynthesis:
    """CUDA-graph replay of ``SynthesisNetwork.forward`` for a fixed batch size (inference): the ~175 launches of a
    256x256 forward are captured once - the tensor maps are kernel parameters, so the graph holds them - and replayed
    with one host call.  ``noise_mode='random'`` draws fresh noise on every replay (torch registers the generator
    state with the graph)."""

    def __init__(self, synthesis, batch, noise_mode="const"):
        p = next(synthesis.parameters())
        self.synthesis, self.noise_mode = synthesis, noise_mode
        self.ws = torch.zeros([batch, synthesis.num_ws, synthesis.w_dim], device=p.device)
        side = torch.cuda.Stream(device=p.device)
        side.wait_stream(torch.cuda.current_stream(p.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):  # packs the frozen weights, binds contexts, fills the allocator
                synthesis(self.ws, noise_mode=noise_mode)
        torch.cuda.current_stream(p.device).wait_stream(side)
        from .. import _lib
        count0 = _lib.load().vcn_launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.img = synthesis(self.ws, noise_mode=noise_mode)
        self.launches = int(_lib.load().vcn_launch_count() - count0)  # libvcn kernels one replay launches

    def __call__(self, ws):
        self.ws.copy_(ws, non_blocking=True)
        self.graph.replay()
        return self.img
