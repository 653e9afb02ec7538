"""psi_fast - the trainable concept extractor Psi (viscoin/models/concept_extractors.py; mirrored by
viscoin_models.ConceptExtractor) on an explicit bf16 path for the VisCoIN train step.

Under ``torch.autocast`` every step re-lays-out and re-casts Psi's 94 M fp32 master weights for cuDNN (fp32 NCHW ->
channels_last -> bf16: two passes per weight), and on the way back casts each bf16 weight gradient to fp32, converts its
layout and ADDS it into the (freshly zeroed) flat gradient bucket: 26 bytes of traffic per parameter for 6 bytes of
information - ~0.6 ms of copy kernels per step on a B200.  Here:

  * ``_ShadowWeight``: master weight -> bf16 (channels_last for 4-D) in ONE pass; its backward hands the bf16 gradient to
    ``GradBuckets.deposit`` which writes it, cast, straight into the bucket slice (the first contribution of a step is a
    copy, not an add) and fires the bucket's all-reduce bookkeeping;
  * the convolutions / linears run on bf16 tensors directly (same arithmetic as autocast: bf16 operands, fp32 accumulation,
    bf16 outputs);
  * ``AdaptiveAvgPool2d(3)`` on the 8x8 maps is a [9 x 64] matrix product (ATen's NHWC kernel launches 16-32 CTAs: 40 us).

conv5.weight also feeds the orthogonality loss in fp32 (losses.py:83-96): it stays on the ordinary autograd path.
"""

import math

import torch
import torch.nn.functional as F


class _ShadowWeight(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p, buckets):
        ctx.p, ctx.buckets = p, buckets
        q = p.detach()
        if q.ndim == 4:
            return q.to(dtype=torch.bfloat16, memory_format=torch.channels_last)
        return q.to(torch.bfloat16)

    @staticmethod
    def backward(ctx, g):
        ctx.buckets.deposit(ctx.p, g)
        return None, None


_pool_mats = {}


def _pool3_matrix(h, w, device):
    """AdaptiveAvgPool2d(3) as a matrix: bin i covers [floor(i*h/3), ceil((i+1)*h/3)).  Returns ([h*w, 9] bf16 0/1, 1/size)."""
    key = (h, w, str(device))
    if key not in _pool_mats:
        m = torch.zeros(h * w, 9)
        for i in range(3):
            y0, y1 = (i * h) // 3, -((-(i + 1) * h) // 3)
            for j in range(3):
                x0, x1 = (j * w) // 3, -((-(j + 1) * w) // 3)
                for y in range(y0, y1):
                    for x in range(x0, x1):
                        m[y * w + x, i * 3 + j] = 1.0
        # 0/1 membership (exact in bf16) and the fp32 reciprocal bin sizes: 1/9 or 1/12 as bf16 GEMM weights would bias every
        # bin by up to 0.2 %
        _pool_mats[key] = (m.to(device=device, dtype=torch.bfloat16), (1.0 / m.sum(dim=0)).to(device).view(1, 9, 1))
    return _pool_mats[key]


def _pool3(x):
    """AdaptiveAvgPool2d(3) of a bf16 channels_last [N,C,H,W] tensor -> [N,C,3,3] (logical NCHW), as one batched GEMM."""
    n, c, h, w = x.shape
    xs = x.permute(0, 2, 3, 1).reshape(n, h * w, c)           # NHWC memory: a view
    p, inv = _pool3_matrix(h, w, x.device)                     # [hw, 9] membership, [1, 9, 1] fp32 1 / bin size
    out = (torch.matmul(p.t().unsqueeze(0), xs).float() * inv).to(x.dtype)  # [n, 9, c]: bin sums (fp32 accumulation) -> means
    return out.permute(0, 2, 1).reshape(n, c, 3, 3)


def supported(psi, hidden):
    return (all(t.is_cuda and t.dtype == torch.bfloat16 for t in hidden) and hasattr(psi, "conv5") and hasattr(psi, "linear2")
            and isinstance(psi.pool, torch.nn.AdaptiveAvgPool2d) and psi.pool.output_size in (3, (3, 3)))


def forward(psi, hidden, buckets):
    """ConceptExtractor.forward (viscoin_models.py / concept_extractors.py) on bf16 tensors.  Returns (concepts, extra) in bf16."""
    W = lambda p: _ShadowWeight.apply(p, buckets)
    b16 = lambda b: None if b is None else b.to(torch.bfloat16)
    conv = lambda x, m, w: F.conv2d(x, w, b16(m.bias), m.stride, m.padding)
    h0, h1, h2 = [t.contiguous(memory_format=torch.channels_last) for t in hidden]
    x1 = F.relu(conv(h0, psi.conv1, W(psi.conv1.weight)))
    x2 = F.relu(conv(h1, psi.conv2, W(psi.conv2.weight)))
    x3 = F.relu(conv(h2, psi.conv3, W(psi.conv3.weight)))
    x = F.relu(conv(torch.cat((x1, x2, x3), 1), psi.conv4, W(psi.conv4.weight)))
    # conv5.weight is also read in fp32 by the orthogonality loss: ordinary autograd path (a second, direct deposit would race it)
    w5 = psi.conv5.weight.to(dtype=torch.bfloat16, memory_format=torch.channels_last)
    c = _pool3(F.relu(conv(x, psi.conv5, w5)))
    e = _pool3(x).reshape(-1, psi.latent_channels * 3 * 9)
    e = F.relu(F.linear(e, W(psi.linear1.weight), b16(psi.linear1.bias)))
    e = F.relu(F.linear(e, W(psi.linear2.weight), b16(psi.linear2.bias)))
    return c, e
