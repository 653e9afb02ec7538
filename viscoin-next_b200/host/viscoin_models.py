"""viscoin_models - the networks around the synthesis path in one VisCoIN training iteration, with the reference's
parameter names (their ``state_dict``s load unchanged).  These are NOT hot-path kernels: plain PyTorch modules over
library convolutions (cuDNN) - the hot path they feed is ``networks.GeneratorAdapted`` / ``networks.Generator``.

  Classifier        viscoin/models/classifiers.py:13-92   (HF ``ResNetModel(ResNetConfig())`` = ResNet-50 + Linear)
  ConceptExtractor  viscoin/models/concept_extractors.py:20-141   (Psi)
  Explainer         viscoin/models/explainers.py:13-56            (Theta)
  LPIPS             third-party ``lpips`` v0.1.4 (flake.nix:33-42), net='vgg', version 0.1: restated from the published
                    algorithm on torchvision's VGG-16 feature stack (the package is not installable offline; random-init)
"""

import torch
from torch import nn


class Classifier(nn.Module):
    def __init__(self, output_classes=200):
        super().__init__()
        from transformers import ResNetConfig, ResNetModel  # third-party arithmetic, same as the reference
        self.resnet = ResNetModel(ResNetConfig())
        self.last_size = self.resnet.config.hidden_sizes[-1]
        self.linear = nn.Linear(self.last_size, output_classes)
        nn.init.kaiming_normal_(self.linear.weight)

    def forward(self, x):
        results = self.resnet(x, output_hidden_states=True)
        output_state = results["pooler_output"].view(-1, self.last_size)
        return self.linear(output_state), results["hidden_states"]


def condition_random_init(model, residual_gain=0.25):
    """Random-init harness only.  HF initialises every BatchNorm weight to 1, so each bottleneck adds a unit-variance
    branch to its shortcut and a random-init ResNet-50 is a chaotic map: measured on a B200, a relative input perturbation
    grows ~1.3x per bottleneck - 16 blocks turn bf16's 4e-3 rounding into a 50 % error of the last hidden state whatever
    the implementation (cuDNN under autocast and the fused path alike), which buries the arithmetic a parity test is
    meant to compare.  Trained networks are not like that; the standard remedy at initialisation ("zero-init residual",
    Goyal et al. 2017) scales the LAST BatchNorm weight of every residual branch down.  Same FLOPs, same kernels."""
    n = 0
    with torch.no_grad():
        for m in model.modules():
            layer = getattr(m, "layer", None)
            if isinstance(layer, nn.Sequential) and hasattr(m, "shortcut") and len(layer) > 0:
                bn = getattr(layer[-1], "normalization", None)
                if isinstance(bn, nn.BatchNorm2d):
                    bn.weight.mul_(residual_gain)
                    n += 1
    return n


def calibrate_batchnorm(model, images):
    """Random-init harness only (the reference always loads a trained f): one forward in train mode with cumulative
    averaging so that every BatchNorm's running statistics equal the batch statistics of `images`.  Without it a
    random-init ResNet-50 in eval mode (running mean 0 / var 1) produces logits of magnitude 1e2..1e3, for which the
    reference's own ``softmax(...).log()`` (losses.py:27) is -inf * 0 = NaN."""
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.reset_running_stats()
        m.momentum = None
    was_training = model.training
    model.train()
    with torch.no_grad():
        model(images)
    model.train(was_training)
    for m in bns:
        m.momentum = 0.1


def fold_batchnorm(model):
    """Frozen, eval-mode classifier only (viscoin.py:114-115): every `conv -> BatchNorm2d` pair of the HF ResNet
    (`ResNetConvLayer.convolution / .normalization`) becomes one convolution with the affine map folded into its weight and
    a bias; the BatchNorm is replaced by Identity.  Same function (up to fp rounding), ~100 fewer kernels per pass."""
    n = 0
    for m in model.modules():
        conv, bn = getattr(m, "convolution", None), getattr(m, "normalization", None)
        if isinstance(conv, nn.Conv2d) and isinstance(bn, nn.BatchNorm2d):
            with torch.no_grad():
                scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
                w = conv.weight * scale.reshape(-1, 1, 1, 1)
                b = bn.bias - bn.running_mean * scale
                if conv.bias is not None:
                    b = b + conv.bias * scale
                fused = nn.Conv2d(conv.in_channels, conv.out_channels, conv.kernel_size, conv.stride, conv.padding,
                                  conv.dilation, conv.groups, bias=True).to(device=w.device, dtype=w.dtype)
                fused.weight.copy_(w)
                fused.bias.copy_(b)
                fused.requires_grad_(False)
            m.convolution = fused
            m.normalization = nn.Identity()
            n += 1
    return n


class ConvBiasAct(nn.Conv2d):
    """nn.Conv2d whose bias add and ReLU run as ONE ``vcn_bias_act`` pass over the convolution output (cuDNN adds the bias
    with a separate elementwise kernel and the ReLU is a third one: two extra read+write passes over every activation of
    the frozen ResNet-50 / VGG-16).  Same parameters and ``state_dict`` names as the Conv2d it replaces (the instance's
    class is swapped in place by ``fuse_conv_bias_relu``); CPU tensors take the plain torch path."""
    vcn_act = "relu"

    def forward(self, x):
        if not x.is_cuda or self.bias is None or self.padding_mode != "zeros":
            y = super().forward(x)
            return torch.relu(y) if self.vcn_act == "relu" else y
        from . import bias_act as _bias_act
        y = torch.nn.functional.conv2d(x, self.weight, None, self.stride, self.padding, self.dilation, self.groups)
        return _bias_act.bias_act(y, self.bias.to(y.dtype), act=self.vcn_act)


def fuse_conv_bias_relu(model):
    """Frozen eval-mode networks only.  torchvision VGG (`Conv2d, ReLU` neighbours in a Sequential) and the HF ResNet after
    ``fold_batchnorm`` (`ResNetConvLayer.convolution / .activation`): the Conv2d becomes a ``ConvBiasAct`` and the ReLU module
    an Identity.  Returns the number of fused pairs."""
    n = 0
    for m in model.modules():
        if isinstance(m, nn.Sequential):
            names = list(m._modules.keys())
            for a, b in zip(names, names[1:]):
                ca, rb = m._modules[a], m._modules[b]
                if type(ca) is nn.Conv2d and ca.bias is not None and isinstance(rb, nn.ReLU):
                    ca.__class__ = ConvBiasAct
                    m._modules[b] = nn.Identity()
                    n += 1
        conv, act = getattr(m, "convolution", None), getattr(m, "activation", None)
        if type(conv) is nn.Conv2d and conv.bias is not None and isinstance(act, nn.ReLU) \
                and isinstance(getattr(m, "normalization", None), nn.Identity):
            conv.__class__ = ConvBiasAct
            m.activation = nn.Identity()
            n += 1
    return n


class ConceptExtractor(nn.Module):
    def __init__(self, n_concepts=256, latent_channels=512, input_channels1=512, input_channels2=1024, input_channels3=2048):
        super().__init__()
        self.latent_channels, self.n_concepts = latent_channels, n_concepts
        self.conv1 = nn.Conv2d(input_channels1, latent_channels, kernel_size=5, padding=2, stride=4)
        self.conv2 = nn.Conv2d(input_channels2, latent_channels, kernel_size=3, padding=1, stride=2)
        self.conv3 = nn.Conv2d(input_channels3, latent_channels, kernel_size=3, padding=1, stride=1)
        self.conv4 = nn.Conv2d(3 * latent_channels, 3 * latent_channels, kernel_size=3, padding=1, stride=1)
        self.pool = nn.AdaptiveAvgPool2d(3)
        self.conv5 = nn.Conv2d(3 * latent_channels, n_concepts, kernel_size=1, padding=0, stride=1)
        self.linear1 = nn.Linear(latent_channels * 3 * 9, 9 * n_concepts, bias=True)
        self.linear2 = nn.Linear(9 * n_concepts, 9 * n_concepts, bias=False)
        self.activ = nn.ReLU()

    def forward(self, latent_layers):
        assert len(latent_layers) == 3, "Input should be a tuple of 3 tensors"
        x1 = self.activ(self.conv1(latent_layers[0]))
        x2 = self.activ(self.conv2(latent_layers[1]))
        x3 = self.activ(self.conv3(latent_layers[2]))
        x = self.activ(self.conv4(torch.cat((x1, x2, x3), 1)))
        x1 = self.pool(self.activ(self.conv5(x)))
        x2 = self.pool(x).reshape(-1, self.latent_channels * 3 * 9)
        x2 = self.activ(self.linear1(x2))
        x2 = self.activ(self.linear2(x2))
        return x1, x2


class Explainer(nn.Module):
    def __init__(self, n_concepts=256, n_classes=200):
        super().__init__()
        self.dropout = nn.Dropout(0.01)
        self.pool = nn.AdaptiveMaxPool2d(1)
        self.linear = nn.Linear(n_concepts, n_classes, bias=True)

    def forward(self, x):
        return self.linear(self.pool(self.dropout(x)).flatten(start_dim=1))


_VGG_SLICES = [(0, 4), (4, 9), (9, 16), (16, 23), (23, 30)]  # relu1_2, relu2_2, relu3_3, relu4_3, relu5_3
_VGG_CHANNELS = [64, 128, 256, 512, 512]


class LPIPS(nn.Module):
    """lpips.LPIPS(net='vgg') forward: scale -> 5 VGG-16 feature taps -> unit-normalise over channels -> squared
    difference -> non-negative 1x1 'lin' weights -> spatial mean -> sum over taps.  Output [N,1,1,1]."""

    def __init__(self):
        super().__init__()
        import torchvision
        feats = torchvision.models.vgg16(weights=None).features
        self.slices = nn.ModuleList([nn.Sequential(*[feats[i] for i in range(a, b)]) for a, b in _VGG_SLICES])
        self.lins = nn.ParameterList([nn.Parameter(torch.rand(1, c, 1, 1) / c) for c in _VGG_CHANNELS])
        self.register_buffer("shift", torch.tensor([-0.030, -0.088, -0.188]).view(1, 3, 1, 1))
        self.register_buffer("scale", torch.tensor([0.458, 0.448, 0.450]).view(1, 3, 1, 1))
        self.requires_grad_(False)
        self.eval()
        self.fused = True  # CUDA tensors: fused tap kernels (False = plain torch formulation)
        self.feature_fn = None  # frozen_nets.FusedVGGFeatures: the bf16 channels_last fast path of the feature stack

    def features(self, x):
        if self.feature_fn is not None and x.is_cuda:
            return self.feature_fn(x)
        out = []
        x = (x - self.shift) / self.scale
        for s in self.slices:
            x = s(x)
            out.append(x)
        return out

    @staticmethod
    def _unit(f, eps=1e-10):
        # lpips.normalize_tensor: f / (sqrt(sum f^2) + eps).  The 1e-30 under the root only matters where a pixel's
        # feature vector is exactly zero: the value is unchanged (0 / eps) but the gradient is 0 instead of 0 * inf = NaN
        f = f.float()
        return f / ((f.square().sum(dim=1, keepdim=True) + 1e-30).sqrt() + eps)

    def forward(self, in0, in1, target_features=None):
        """in0: the image that may carry a gradient (rebuilt), in1: the target.  On a CUDA device each tap's distance is
        ONE fused kernel forward and one backward (lpips_tap) instead of ~20 full-size fp32 elementwise passes; the
        target's features are computed without a graph (or passed in: `target_features = self.features(in1)` computed
        earlier, e.g. on a side stream - they depend on nothing but the target images)."""
        if in0.is_cuda and self.fused:
            from . import lpips_tap as _tap
            if target_features is not None:
                f1 = target_features
            else:
                with torch.no_grad():
                    f1 = self.features(in1)
            f0 = self.features(in0)
            val = 0
            for k in range(len(f0)):
                val = val + _tap.lpips_tap(f0[k], f1[k], self.lins[k])
            return val.reshape(-1, 1, 1, 1)
        n = in0.shape[0]
        f = self.features(torch.cat([in0, in1]))  # one batched pass over both image sets
        val = 0
        for k, fk in enumerate(f):
            d = (self._unit(fk[:n]) - self._unit(fk[n:])).square()
            val = val + (d * self.lins[k]).sum(dim=1, keepdim=True).mean(dim=[2, 3], keepdim=True)
        return val
