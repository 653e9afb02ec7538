"""psi_fast (the trainable concept extractor on an explicit bf16 path with weight shadows and direct gradient deposit,
viscoin_next_b200/host/psi_fast.py) against the bf16-autocast path it replaces (viscoin/models/concept_extractors.py through
viscoin_models.ConceptExtractor): same arithmetic class (bf16 operands and storage, fp32 accumulation), so outputs and every
parameter gradient must agree to bf16 noise: rel-L2 <= 2e-2, cosine >= 0.999; the bucket bookkeeping must see every parameter."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel2(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30)), float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-30))


def test_psi_fast_matches_autocast_path():
    import viscoin_next_b200 as V
    from viscoin_next_b200.host import psi_fast
    dev = torch.device("cuda:0")
    torch.manual_seed(4)
    psi = V.viscoin_models.ConceptExtractor().to(dev)
    buckets = V.dp.GradBuckets(list(psi.parameters()))
    g = torch.Generator(device=dev).manual_seed(1)
    cl = lambda t: t.contiguous(memory_format=torch.channels_last)
    hidden = [cl(torch.relu(torch.randn(2, c, r, r, device=dev, generator=g)).bfloat16()) for c, r in ((512, 32), (1024, 16), (2048, 8))]
    assert psi_fast.supported(psi, hidden)

    def loss_of(c, e):
        return c.float().square().mean() + e.float().square().mean() + V.viscoin_step.concept_orthogonality_loss(psi)

    def grads():
        out = {k: p.grad.detach().clone() for k, p in psi.named_parameters()}
        assert all(v == 0 for v in buckets._pending), buckets._pending  # every parameter arrived
        buckets.reset()
        return out

    c1, e1 = psi_fast.forward(psi, hidden, buckets)
    loss_of(c1, e1).backward()
    fast = grads()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        c0, e0 = psi(hidden)
    loss_of(c0, e0).backward()
    ref = grads()
    for name, a, b in (("concepts", c1, c0), ("extra", e1, e0)):
        r, cos = _rel2(a, b)
        print("%-10s rel-L2 %.2e cos %.6f" % (name, r, cos))
        assert a.shape == b.shape and r <= 2e-2 and cos >= 0.999
    for k in ref:
        r, cos = _rel2(fast[k], ref[k])
        print("%-16s rel-L2 %.2e cos %.6f" % (k, r, cos))
        assert r <= 2e-2 and cos >= 0.999, k
    # a second contribution in the same step is ADDED (only the first one may overwrite the zeroed slice)
    w = psi.linear2.weight
    g1 = torch.randn_like(w).bfloat16()
    buckets.deposit(w, g1)
    buckets.deposit(w, g1)
    assert torch.allclose(w.grad, 2 * g1.float())
    buckets.reset()
