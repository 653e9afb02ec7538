"""Full VisCoIN train step on the B200 vs the CPU oracle (oracle/ref_step.py) on the same weights and inputs
(viscoin/training/viscoin.py:140-223): every loss term, every trainable gradient, and the parameters after both Adam
updates - for the fp32-library configuration AND for exactly the configuration ``bench.py`` times (bf16 autocast,
channels_last, folded BatchNorm, bf16-resident synthesis activations, whole-step CUDA graph).

Theta's dropout is off and noise_mode='const' with noise_strength 0.1 (device RNG streams differ from the CPU's); the
networks fix the resolution at 256x256.  The synthesis path computes on bf16 tensor-core operands with fp32 accumulation:
tolerances are the bf16 gates of SURVEY 8(c).  Measured errors go to gpurun_out/parity_r2.json (-> profiles/)."""

import os
import subprocess
import sys

import pytest
import torch

from step_parity import ROOT, compare, make_inputs, metrics, record, set_noise_strength, summarize

pytestmark = pytest.mark.gpu

FP32_CONFIG = dict(amp=None, channels_last=False, fold_bn=False)  # library networks in fp32 (TF32 off below)
BENCH_CONFIG = dict()                                             # VisCoINTrainer defaults = what bench.py builds


class _tf32_off:
    def __enter__(self):
        self.old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False

    def __exit__(self, *exc):
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = self.old


def _big(out):
    """Per-tensor entries only (the '(all N tensors)' group rows are checked separately)."""
    return {k: v for k, v in out["grads"].items() if "(all" not in k}


def _groups(out):
    return {k: v for k, v in out["grads"].items() if "(all" in k}


# ---- fp32 library networks (TF32 off): isolates the synthesis kernels' bf16 operands ----------------------------------


def test_train_step_matches_oracle_l1_l2_reconstruction():
    """beta = gamma = 0: the reconstruction term is L1 + L2 only, so the gradient that enters the synthesis backward is
    smooth in the rebuilt image.  Loss values <= 2e-2 (measured <= 1.2e-3); EVERY trainable gradient tensor (through the
    synthesis dgrad / dstyles kernels, the adapted mapping and Psi) within the bf16 gate rel-Linf <= 5e-2; Adam update
    direction of each parameter group cosine >= 0.9."""
    with _tf32_off():
        out = compare(FP32_CONFIG, 0.0, 0.0, n=1, label="fp32lib_l1l2")
    summarize(out)
    assert all(v["rel"] <= 2e-2 for v in out["losses"].values()), out["losses"]
    bad = {k: v for k, v in _big(out).items() if v["rel_linf"] > 5e-2}
    assert not bad, bad
    assert all(v["cos"] >= 0.999 for v in _groups(out).values()), _groups(out)
    assert all(v["update_cos"] >= 0.9 for v in out["adam"].values()), out["adam"]


def test_train_step_matches_oracle_all_losses():
    """All six terms (beta = 3 LPIPS, gamma = 0.1 classifier agreement): loss values <= 2e-2 (measured <= 2e-3), every
    parameter group's gradient rel-L2 <= 3e-2 and cosine >= 0.9995 (measured 1.4e-2 / 0.99991), Adam update cosine >= 0.98.
    (Round 1 gated this test at cosine 0.8 and blamed the random-init LPIPS: the cause was the random-init ResNet-50 f,
    which amplifies any rounding ~1.3x per block; with the harness' conditioned initialisation - viscoin_models.
    condition_random_init - the bf16-operand synthesis path matches the fp32 oracle as closely here as in the L1+L2 test.)"""
    with _tf32_off():
        out = compare(FP32_CONFIG, 3.0, 0.1, n=1, label="fp32lib_all")
    summarize(out)
    assert all(v["rel"] <= 2e-2 for v in out["losses"].values()), out["losses"]
    bad = {k: v for k, v in _groups(out).items() if v["cos"] < 0.9995 or v["rel_l2"] > 3e-2}
    assert not bad, bad
    assert all(v["update_cos"] >= 0.98 for v in out["adam"].values()), out["adam"]


# ---- the benchmarked configuration ------------------------------------------------------------------------------------


# Gates of the benchmarked configuration = the errors measured on a B200 (profiles/parity_r2.json) with a ~2x margin on
# (1 - cosine) and on rel-L2.  Measured: Psi rel-L2 1.35e-1 / cos 0.9908, adapted mapping 3.2e-2 / 0.99955, Theta 7.8e-3;
# Adam-update cosine 0.893 / 0.979 / 0.996.  test_bench_config_error_decomposition attributes them: the bf16-resident
# synthesis kernels alone give 1.0e-2 (cos 0.99995) - the rest is bf16 autocast arithmetic in the LIBRARY networks f / Psi
# (plain torch autocast alone: 1.26e-1 / 3.7e-2).  Yardstick on the same harness: the reference's own modules with torch's
# default TF32 convolutions against its CPU path: Psi 4.0e-2, mapping 5.9e-3 (tf32_reference_modules_all).
BENCH_GATES = {  # group -> (min cosine, max rel-L2, min Adam-update cosine)
    "concept_extractor": (0.98, 0.27, 0.79),
    "explainer": (0.999, 0.03, 0.98),
    "viscoin_gan": (0.999, 0.065, 0.955),
}


def _check_bench(out, all_losses):
    assert all(v["rel"] <= 1e-2 for v in out["losses"].values()), out["losses"]  # measured <= 1.5e-3
    for name, v in _groups(out).items():
        cos, l2, upd = BENCH_GATES[name.split(" ")[0]]
        assert v["cos"] >= cos and v["rel_l2"] <= l2, (name, v)
    for name, v in out["adam"].items():
        assert v["update_cos"] >= BENCH_GATES[name][2] and v["param_max_abs_diff_over_lr"] <= 2.5, (name, v)


def test_bench_config_matches_oracle_l1_l2_reconstruction():
    """VisCoINTrainer defaults (bf16 autocast for f / Psi / VGG, channels_last, folded BN, bf16-resident synthesis), two
    real + two fake images, L1+L2 reconstruction: loss values (<= 3e-2, measured <= 1.4e-2), all gradients, parameters
    after both Adam updates (Adam's first step is lr * sign(g) for |g| >> eps: |p - p_ref| <= 2 lr by construction, the
    informative numbers are the update cosine and the sign agreement)."""
    out = compare(BENCH_CONFIG, 0.0, 0.0, n=2, label="bench_l1l2")
    summarize(out)
    _check_bench(out, False)


def test_bench_config_matches_oracle_all_losses():
    """Same with all six loss terms."""
    out = compare(BENCH_CONFIG, 3.0, 0.1, n=2, label="bench_all")
    summarize(out)
    _check_bench(out, True)


def test_bench_config_error_decomposition():
    """Where the bench configuration's gradient error comes from (L1+L2 reconstruction, same inputs): one source switched
    on at a time.  Recorded in the parity report; gated only loosely (each single source must stay below the full bench
    configuration's gates)."""
    rows = {
        "resident_synthesis_only": dict(amp=None, channels_last=False, fold_bn=False, resident_bf16=True),
        "bf16_autocast_only": dict(resident_bf16=False),
        "bf16_f_vgg_fp32_psi": dict(amp_psi=None),
    }
    for label, cfg in rows.items():
        if cfg.get("amp", "bf16") is None:
            with _tf32_off():
                out = compare(cfg, 0.0, 0.0, n=2, label="decomp_" + label, with_adam=False)
        else:
            out = compare(cfg, 0.0, 0.0, n=2, label="decomp_" + label, with_adam=False)
        print("==", label)
        summarize(out)
        for name, v in _groups(out).items():
            cos, l2, _ = BENCH_GATES[name.split(" ")[0]]
            assert v["cos"] >= cos and v["rel_l2"] <= l2, (label, name, v)


def test_step_graphed_equals_eager_step():
    """``step_graphed`` (what bench.py times: the whole iteration replayed from one CUDA graph) against the eager ``step``
    on an identical replica: same z stream, noise_mode='const', dropout off, four iterations (two eager warm-ups, the
    capture, one replay).  The yardstick is a SECOND eager replica: a few reductions of the step accumulate with atomics
    (bias / style gradient sums, cuDNN's split-k weight gradients), so two eager runs already differ in the last bits and
    Adam's first steps (lr * sign(g) per element) turn that into +-lr flips of near-zero-gradient elements.  The graphed
    run must sit inside that run-to-run noise (3x margin), i.e. capturing the step does not change what it computes."""
    import viscoin_next_b200 as V
    dev = torch.device("cuda:0")
    trs = []
    for _ in range(3):
        # noise strengths stay at their initial 0: `synthetic_samples` draws its per-layer noise from the device RNG
        # (utils.py:37-38 passes no noise_mode), whose stream a captured graph and an eager run do not share
        tr = V.viscoin_step.VisCoINTrainer(dev, seed=3, deterministic=True)
        tr.synthesis_kwargs = {"noise_mode": "const"}
        trs.append(tr)
    eager, eager2, graphed = trs
    g = torch.Generator().manual_seed(0)
    real = torch.randn(2, 3, 256, 256, generator=g)
    labels = torch.randint(0, 200, (2,), generator=g)
    gens = [torch.Generator().manual_seed(9) for _ in range(3)]
    worst = {"graphed": 0.0, "eager2": 0.0}
    for it in range(4):
        a = {k: float(v) for k, v in eager.step(real.to(dev), labels.to(dev), 1000, z_generator=gens[0]).items()}
        b = {k: float(v) for k, v in eager2.step(real.to(dev), labels.to(dev), 1000, z_generator=gens[1]).items()}
        c = {k: float(v) for k, v in graphed.step_graphed(real, labels, 1000, z_generator=gens[2]).items()}
        for k in a:
            worst["eager2"] = max(worst["eager2"], abs(a[k] - b[k]) / max(abs(a[k]), 1e-6))
            worst["graphed"] = max(worst["graphed"], abs(a[k] - c[k]) / max(abs(a[k]), 1e-6))
        print("iteration %d: %s" % (it, {k: "%.3e / %.3e" % (abs(a[k] - b[k]), abs(a[k] - c[k])) for k in a}))
    assert (2, True) in graphed._graphs, "the graph was never captured"
    flat = lambda tr: torch.cat([p.detach().flatten() for p in tr.concept_extractor.parameters()] +
                                [p.detach().flatten() for p in tr.viscoin_gan.mapping.parameters()])
    pe, pe2, pg = flat(eager), flat(eager2), flat(graphed)
    m2, mg = metrics(pe2, pe), metrics(pg, pe)
    print("eager2 vs eager : worst loss rel diff %.2e, params rel-L2 %.2e max-abs %.2e" % (worst["eager2"], m2["rel_l2"], m2["max_abs"]))
    print("graphed vs eager: worst loss rel diff %.2e, params rel-L2 %.2e max-abs %.2e" % (worst["graphed"], mg["rel_l2"], mg["max_abs"]))
    record("graphed_vs_eager", {"worst_loss_rel": worst["graphed"], "params": mg, "bitwise_equal": bool(torch.equal(pe, pg)), "iterations": 4,
                                "eager_rerun_worst_loss_rel": worst["eager2"], "eager_rerun_params": m2})
    assert worst["graphed"] <= max(3 * worst["eager2"], 1e-5) and mg["rel_l2"] <= max(3 * m2["rel_l2"], 1e-7)


def test_all_losses_fp32_reference_modules_conditioning():
    """Settles where the loose all-loss gradient gate comes from: the SAME harness with the synthesis
    path run by the reference's own modules in fp32 on the GPU (its plugins / torch ops + cuDNN with TF32 off, no kernel of
    this repo) against the same CPU oracle.  If this reaches cosine ~1 the 0.88..0.98 of the product path is bf16 operand
    rounding amplified by the random-init perceptual loss; the measured values are recorded either way."""
    from oracle import stage_ref
    if not stage_ref.available():
        pytest.skip("oracle/_ref not staged (build container only stages it when /root/reference is mounted)")
    import viscoin_next_b200 as V
    try:
        ref = stage_ref.load_plugins()
    except Exception as e:  # the reference itself would fall back to its torch ops here
        print("reference plugins unavailable (%s): reference runs its impl='ref' path on the GPU" % e)
        ref = stage_ref.import_reference()
    torch.manual_seed(11)
    VM, N = V.viscoin_models, V.networks
    mine = (VM.Classifier(200), VM.ConceptExtractor(), VM.Explainer(n_classes=200),
            N.Generator(z_dim=512, c_dim=0, w_dim=512, img_resolution=256, img_channels=3, mapping_kwargs={"num_layers": 2}))
    lp = VM.LPIPS()
    lp.fused = False
    G = ref.networks.Generator(z_dim=512, c_dim=0, w_dim=512, img_resolution=256, img_channels=3, mapping_kwargs={"num_layers": 2})
    G.load_state_dict(mine[3].state_dict(), strict=True)
    GA = ref.gan.GeneratorAdapted.from_gan(G)
    GA.mapping.load_state_dict(N.GeneratorAdapted.from_gan(mine[3]).mapping.state_dict(), strict=True)
    psi = ref.concept_extractors.ConceptExtractor()
    psi.load_state_dict(mine[1].state_dict(), strict=True)
    theta = ref.explainers.Explainer(n_classes=200)
    theta.load_state_dict(mine[2].state_dict(), strict=True)
    # BatchNorm statistics of the random-init classifier (what VisCoINTrainer does when it builds the models itself)
    cls = mine[0].cuda()
    VM.condition_random_init(cls)  # as VisCoINTrainer does for its own random-init harness
    with torch.no_grad():
        fakes = G.cuda()(torch.randn(4, 512, generator=torch.Generator().manual_seed(11)).cuda(), None).float()
    VM.calibrate_batchnorm(cls, torch.cat([torch.randn(4, 3, 256, 256, generator=torch.Generator().manual_seed(12)).cuda(), fakes]))
    with _tf32_off():
        out = compare(FP32_CONFIG, 3.0, 0.1, n=1, label="fp32_reference_modules_all", with_adam=False,
                      models=(cls, psi, theta, G, GA, lp))
    summarize(out)
    assert all(v["rel"] <= 1e-3 for v in out["losses"].values()), out["losses"]
    assert all(v["cos"] >= 0.8 for v in _groups(out).values()), _groups(out)
    # the yardstick: the reference's DEFAULT GPU arithmetic (torch leaves cudnn.allow_tf32 on; only the GAN trainer turns it
    # off, training_loop.py:128-129) against its own CPU path on this harness
    out = compare(FP32_CONFIG, 3.0, 0.1, n=1, label="tf32_reference_modules_all", with_adam=False,
                  models=(cls, psi, theta, G, GA, lp))
    summarize(out)


def test_train_step_updates_and_is_finite():
    """bf16-autocast configuration of the bench: two steps run, losses finite, trainable weights move, frozen ones do not."""
    import viscoin_next_b200 as V
    dev = torch.device("cuda:0")
    tr = V.viscoin_step.VisCoINTrainer(dev, seed=3)
    g = torch.Generator(device=dev).manual_seed(0)
    real = torch.randn(2, 3, 256, 256, device=dev, generator=g)
    labels = torch.randint(0, 200, (2,), device=dev, generator=g)
    w_map = tr.viscoin_gan.mapping.s3[0].weight.detach().clone()
    w_psi = tr.concept_extractor.conv4.weight.detach().clone()
    w_syn = tr.viscoin_gan.synthesis.b64.conv1.weight.detach().clone()
    w_f = tr.classifier.linear.weight.detach().clone()
    for i in range(2):
        out = tr.step(real, labels, 1000)
        assert all(bool(torch.isfinite(v)) for v in out.values()), out
    assert not torch.equal(w_map, tr.viscoin_gan.mapping.s3[0].weight)
    assert not torch.equal(w_psi, tr.concept_extractor.conv4.weight)
    assert torch.equal(w_syn, tr.viscoin_gan.synthesis.b64.conv1.weight)
    assert torch.equal(w_f, tr.classifier.linear.weight)
    assert tr.generator_gan.synthesis is tr.viscoin_gan.synthesis


def test_data_parallel_two_ranks_equal_one_rank_on_concatenated_batch():
    """SURVEY App. E 'DP' row on hardware: the real step on 2 NCCL ranks (each its own shard, bucketed all-reduce
    overlapped with backward) gives the gradients of ONE rank on the concatenated batch.  Needs 2 GPUs (the round-end run
    has one: run with `gpurun --gpus 2`; the measured errors are committed in profiles/parity_r2.json)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "dp_check.py")]
    proc = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    print(proc.stdout[-4000:])
    assert proc.returncode == 0, "dp_check.py failed"
