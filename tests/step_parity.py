"""Shared machinery of the train-step parity tests (not a test module): one VisCoIN iteration on the B200 - in any
trainer configuration, including exactly the one ``bench.py`` times - against the CPU oracle (``oracle/ref_step.py``,
fp32) on the same weights and inputs: loss values, gradients, and parameters after both Adam updates
(viscoin/training/viscoin.py:140-223).  Measured errors are appended to ``gpurun_out/parity_r2.json`` (copied to
``profiles/`` by hand after a GPU run) so that the numbers behind every gate are on record."""

import json
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPORT = os.path.join(ROOT, "gpurun_out", "parity_r2.json")


def record(section, payload):
    """Merge `payload` under `section` into gpurun_out/parity_r2.json (best effort: never fails a test)."""
    try:
        os.makedirs(os.path.dirname(REPORT), exist_ok=True)
        data = {}
        if os.path.exists(REPORT):
            with open(REPORT) as fh:
                data = json.load(fh)
        data[section] = payload
        with open(REPORT, "w") as fh:
            json.dump(data, fh, indent=1, sort_keys=True)
    except Exception as e:  # pragma: no cover
        print("parity report not written:", e)


def metrics(a, ref):
    """rel-Linf (max-abs error over max-abs reference), rel-L2, cosine - all in float64 on the CPU."""
    a, ref = a.detach().double().cpu().flatten(), ref.detach().double().cpu().flatten()
    d = a - ref
    return {"rel_linf": float(d.abs().max() / ref.abs().max().clamp_min(1e-30)),
            "rel_l2": float(d.norm() / ref.norm().clamp_min(1e-300)),
            "cos": float(torch.dot(a, ref) / (a.norm() * ref.norm()).clamp_min(1e-300)),
            "max_abs": float(d.abs().max())}


def set_noise_strength(trainer, value=0.1):
    with torch.no_grad():
        for name, p in trainer.viscoin_gan.synthesis.named_parameters():
            if name.endswith("noise_strength"):
                p.fill_(value)


def make_inputs(trainer, n, seed=5):
    import viscoin_next_b200 as V
    g = torch.Generator().manual_seed(seed)
    real = torch.randn(n, 3, 256, 256, generator=g)
    labels = torch.randint(0, 200, (n,), generator=g)
    with torch.no_grad():
        fake = V.viscoin_step.synthetic_samples(trainer.generator_gan, n, trainer.device, g).float()
    assert fake.shape == (n, 3, 256, 256) and bool(torch.isfinite(fake).all())
    return real, labels, fake


TRAINABLE_GROUPS = ("concept_extractor", "explainer", "viscoin_gan")


def oracle_step(state, real, labels, fake, beta, gamma, lr=1e-4, adam=True):
    """CPU oracle: losses, gradients of EVERY trainable tensor (Psi, Theta, adapted mapping) and - with adam - the
    parameters after one fp32 torch.optim.Adam step on those gradients (viscoin.py:130-134, 215-223)."""
    from oracle import ref_step as RS
    leaves = {}
    for grp in TRAINABLE_GROUPS:
        for k, v in state[grp].items():
            if v.is_floating_point() and (grp != "viscoin_gan" or k.startswith("mapping.")):
                v.requires_grad_(True)
                leaves[grp + "." + k] = v
    want = RS.step_losses(state, real, labels, fake.cpu(), noise_mode="const", beta=beta, gamma=gamma)
    want.pop("_rebuilt")
    sum(want.values()).backward()
    grads = {k: (v.grad.detach().clone() if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items()}
    after = None
    if adam:
        before = {k: v.detach().clone() for k, v in leaves.items()}
        for v in leaves.values():
            if v.grad is None:
                v.grad = torch.zeros_like(v)
        torch.optim.Adam(list(leaves.values()), lr=lr).step()
        after = {k: (v.detach().clone(), before[k]) for k, v in leaves.items()}
    return {k: float(v) for k, v in want.items()}, grads, after


def trainer_named_params(tr):
    out = {}
    for grp, mod in (("concept_extractor", tr.concept_extractor), ("explainer", tr.explainer)):
        for k, p in mod.named_parameters():
            out[grp + "." + k] = p
    for k, p in tr.viscoin_gan.mapping.named_parameters():
        out["viscoin_gan.mapping." + k] = p
    return out


def compare(trainer_kwargs, beta, gamma, n=1, label="", seed=11, with_adam=True, models=None):
    """Build the trainer, run forward+backward (gradients), then - from the same state - one full ``step()`` (both Adam
    updates), and compare everything with the oracle.  Returns the dict of measured errors."""
    import viscoin_next_b200 as V
    dev = torch.device("cuda:0")
    params = V.viscoin_step.VisCoINParams()
    params.beta, params.gamma = beta, gamma
    tr = V.viscoin_step.VisCoINTrainer(dev, seed=seed, deterministic=True, params=params, keep_unfolded=True, models=models,
                                       **trainer_kwargs)
    tr.synthesis_kwargs = {"noise_mode": "const"}
    set_noise_strength(tr, 0.1)
    state = tr.oracle_state()
    real, labels, fake = make_inputs(tr, n)

    losses = tr.forward_losses(real.to(dev), labels.to(dev), 1000, fake_images=fake)
    sum(losses.values()).backward()
    tr.buckets.finish()
    named = trainer_named_params(tr)
    got_grads = {k: p.grad.detach().float().cpu().clone() for k, p in named.items()}
    got_losses = {k: float(v) for k, v in losses.items()}
    tr.buckets.reset()

    want_losses, want_grads, want_after = oracle_step(state, real, labels, fake, beta, gamma, lr=params.learning_rate,
                                                      adam=with_adam)
    out = {"config": {k: str(v) for k, v in trainer_kwargs.items()}, "beta": beta, "gamma": gamma, "n": n, "losses": {},
           "grads": {}, "adam": {}}
    for k, w in want_losses.items():
        out["losses"][k] = {"gpu": got_losses[k], "cpu": w, "rel": abs(got_losses[k] - w) / max(abs(w), 1e-6)}
    # per-tensor gradient errors (the big ones only individually: >= 4096 elements; all of them per group)
    for grp in TRAINABLE_GROUPS:
        keys = [k for k in want_grads if k.startswith(grp + ".")]
        a = torch.cat([got_grads[k].flatten() for k in keys])
        r = torch.cat([want_grads[k].flatten() for k in keys])
        out["grads"][grp + " (all %d tensors)" % len(keys)] = metrics(a, r)
    for k in want_grads:
        if want_grads[k].numel() >= 4096 and float(want_grads[k].abs().max()) > 0:
            out["grads"][k] = metrics(got_grads[k], want_grads[k])

    if with_adam:
        before = {k: p.detach().float().cpu().clone() for k, p in named.items()}
        res = tr.step(real.to(dev), labels.to(dev), 1000, fake_images=fake)  # the SAME state: weights were not touched above
        assert all(bool(torch.isfinite(v)) for v in res.values())
        lr = params.learning_rate
        for grp in TRAINABLE_GROUPS:
            keys = [k for k in want_after if k.startswith(grp + ".")]
            d_gpu = torch.cat([(named[k].detach().float().cpu() - before[k]).flatten() for k in keys]).double()
            d_ref = torch.cat([(want_after[k][0] - want_after[k][1]).flatten() for k in keys]).double()
            p_gpu = torch.cat([named[k].detach().float().cpu().flatten() for k in keys]).double()
            p_ref = torch.cat([want_after[k][0].flatten() for k in keys]).double()
            moved = d_ref.abs() > 0.5 * lr  # elements whose oracle gradient is not ~0 (Adam's first step is lr * g / (|g| + eps))
            out["adam"][grp] = {
                "param_max_abs_diff_over_lr": float((p_gpu - p_ref).abs().max() / lr),
                "param_rel_l2": float((p_gpu - p_ref).norm() / p_ref.norm().clamp_min(1e-300)),
                "update_cos": float(torch.dot(d_gpu, d_ref) / (d_gpu.norm() * d_ref.norm()).clamp_min(1e-300)),
                "update_sign_agree": float((torch.sign(d_gpu[moved]) == torch.sign(d_ref[moved])).double().mean()) if bool(moved.any()) else 1.0,
                "elements": int(d_ref.numel()), "moved": int(moved.sum())}
    record(label or "compare", out)
    return out


def summarize(out):
    for k, v in out["losses"].items():
        print("loss %-6s gpu %.6f cpu %.6f rel %.2e" % (k, v["gpu"], v["cpu"], v["rel"]))
    for k, v in out["grads"].items():
        print("grad %-50s rel-Linf %.2e rel-L2 %.2e cos %.6f" % (k, v["rel_linf"], v["rel_l2"], v["cos"]))
    for k, v in out["adam"].items():
        print("adam %-20s |dp|max/lr %.3f param rel-L2 %.2e update cos %.4f sign-agree %.4f" % (
            k, v["param_max_abs_diff_over_lr"], v["param_rel_l2"], v["update_cos"], v["update_sign_agree"]))
