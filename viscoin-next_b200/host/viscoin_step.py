"""viscoin_step - one VisCoIN training iteration (the consumer of the synthesis hot path) and its
data-parallel form.  Follows viscoin/training/viscoin.py:140-223 statement by statement:

    fake  = synthetic_samples(generator_gan, N)                       utils.py:26-39        hot path, forward only
    classes, hidden = f(cat[real, fake])                              viscoin.py:173
    phi, phi2 = Psi(hidden[-3:]) ; expl = Theta(phi)                  :174-175
    acc / concept-regularisation / output-fidelity / orthogonality    :181-190, losses.py:66-96,165-175
    rebuilt, ws = viscoin_gan(z1=phi, z2=phi2, return_latents=True)   :193-195             hot path, fwd + dx/dstyles bwd
    rec = L1 + L2 + gamma*CE(f(rebuilt), f(x)) + beta*LPIPS           :196-205, losses.py:125-157
    gan = mse(ws, fixed_w_avg)                                        :207, losses.py:178-192
    total.backward() ; Adam(Psi + Theta) ; Adam(viscoin_gan.mapping)  :215-223

The synthesis network is ONE module shared by both generators and frozen (gan.py:142, viscoin.py:116); the classifier
is frozen and in eval mode (:114-115).  Data parallelism = batch sharding only: every rank runs this step on its own
8 real + 8 fake images, and the trainable gradients are averaged by ``dp.GradBuckets`` while backward is still running.
"""

import itertools
import os

import torch
import torch.nn.functional as F

from . import dp as _dp
from . import image_ops as _image_ops
from . import networks as _networks
from . import psi_fast as _psi_fast
from . import viscoin_models as _models


fused_image_ops = os.environ.get("VCN_FUSED_IMAGE_OPS", "1") == "1"


# ---- losses (viscoin/training/losses.py) ------------------------------------------------------------------------------


def cross_cross_entropy_loss(prediction, target_prediction):
    """losses.py:25-29."""
    # log_softmax == softmax(...).log() wherever the latter is finite; it stays finite when a probability underflows
    t = F.softmax(target_prediction.float(), dim=1)
    return (F.log_softmax(prediction.float(), dim=1) * -t).sum(dim=1).mean()


def concept_regularization_loss(concept_embeddings):
    """losses.py:66-80: l1(normalised max-pooled concepts) + l1(concepts)."""
    pooled = F.adaptive_max_pool2d(concept_embeddings, 1).flatten(start_dim=1)
    normed = F.normalize(pooled, p=2, dim=1)
    return normed.abs().mean() + concept_embeddings.abs().mean()


def concept_orthogonality_loss(model):
    """losses.py:83-96."""
    w = model.conv5.weight.view(model.n_concepts, -1)
    nw = F.normalize(w, dim=1).abs()
    return ((nw @ nw.T).sum() - model.n_concepts) / (model.n_concepts ** 2)


def output_fidelity_loss(original_classes, explainer_classes):
    """losses.py:165-175."""
    return cross_cross_entropy_loss(explainer_classes, original_classes.detach())


def reconstruction_loss(reconstructed, original, reconstructed_classes, original_classes, lpips_net, lambda_classes=0.1,
                        lambda_lpips=3.0):
    """losses.py:125-157.  L1 and L2 share one pass over the difference."""
    if fused_image_ops and _image_ops.l1_l2_supported(reconstructed, original):
        l12 = _image_ops.l1_l2(reconstructed, original).sum()  # one pass forward, one backward (csrc/image_ops.cu)
    else:
        diff = reconstructed.float() - original.float()
        l12 = diff.abs().mean() + diff.square().mean()
    return (l12
            + lambda_classes * cross_cross_entropy_loss(reconstructed_classes, original_classes.detach())
            + lambda_lpips * lpips_net(reconstructed, original).mean())


def gan_regularization_loss(gan_latents, model):
    """losses.py:178-192."""
    return (gan_latents - model.mapping.fixed_w_avg.detach()).square().mean()


def synthetic_samples(gan, n_samples, device, generator=None):
    """utils.py:26-39: z ~ N(0,1) on the host, forward through the non-adapted generator."""
    z = torch.randn(n_samples, gan.z_dim, generator=generator)
    return gan(z.to(device), None)


# ---- the step ---------------------------------------------------------------------------------------------------------


class VisCoINParams:
    """VisCoINTrainingParams, viscoin.py:49-81 (CUB defaults)."""
    learning_rate = 1e-4
    cd_fid_iteration = 100
    batch_size = 8
    alpha, beta, gamma, delta = 0.5, 3.0, 0.1, 0.2


class VisCoINTrainer:
    """Holds the five networks + LPIPS + the two Adam optimizers and runs ``step(real, labels, i)``.

    amp: None = fp32 library convolutions (TF32 as torch's cuDNN default, like the reference on a GPU);
         torch.bfloat16 = autocast for the library networks f / Psi / VGG (the synthesis path always runs bf16 tensor-core
         operands with fp32 accumulation)."""

    def __init__(self, device, n_classes=200, seed=0, amp=torch.bfloat16, channels_last=True, params=None, models=None,
                 deterministic=False, fused_adam=None, fold_bn=True, keep_unfolded=False, amp_psi="same", resident_bf16=None,
                 fused_frozen=True, vgg_impl=None):
        self.device = torch.device(device)
        self.p = params or VisCoINParams()
        self.amp = amp if self.device.type == "cuda" else None
        # Psi (trainable) may run at a different precision than the frozen f / VGG-16: "same" = follow `amp`
        self.amp_psi = self.amp if amp_psi == "same" else (amp_psi if self.device.type == "cuda" else None)
        self._resident_override = resident_bf16
        self.psi_fast = os.environ.get("VCN_PSI_FAST", "1") == "1" and self.device.type == "cuda"
        self.channels_last = channels_last and self.device.type == "cuda"
        torch.manual_seed(seed)
        if models is None:
            classifier = _models.Classifier(n_classes)
            concept_extractor = _models.ConceptExtractor()
            explainer = _models.Explainer(n_classes=n_classes)
            generator_gan = _networks.Generator(z_dim=512, c_dim=0, w_dim=512, img_resolution=256, img_channels=3,
                                                mapping_kwargs={"num_layers": 2})
            viscoin_gan = _networks.GeneratorAdapted.from_gan(generator_gan)
            lpips = _models.LPIPS()
        else:
            classifier, concept_extractor, explainer, generator_gan, viscoin_gan, lpips = models
        # viscoin.py:113-123
        classifier.eval().requires_grad_(False)
        viscoin_gan.synthesis.requires_grad_(False)
        generator_gan.requires_grad_(False)  # a trained G_ema snapshot is stored frozen (SURVEY 3.1 footnote)
        self.classifier = classifier.to(self.device)
        self.concept_extractor = concept_extractor.to(self.device)
        self.explainer = explainer.to(self.device)
        self.viscoin_gan = viscoin_gan.to(self.device)
        self.generator_gan = generator_gan.to(self.device)
        self.lpips = lpips.to(self.device)
        self.fold_bn = fold_bn
        if self.channels_last:
            for m in (self.classifier, self.lpips):  # frozen nets only: trainable grads are views into flat buckets
                m.to(memory_format=torch.channels_last)
        if models is None:
            # random-init harness: damp f's residual branches (a random-init ResNet-50 amplifies rounding ~1.3x per block,
            # see condition_random_init) and give its BatchNorms the statistics of the images it will see (noise-like
            # "real" images and smooth generated ones); a trained f ships with its own weights and statistics
            _models.condition_random_init(self.classifier)
            with torch.no_grad():
                fakes = synthetic_samples(self.generator_gan, 4, self.device, torch.Generator().manual_seed(seed)).float()
            noise = torch.randn(4, 3, 256, 256, generator=torch.Generator().manual_seed(seed + 1)).to(self.device)
            _models.calibrate_batchnorm(self.classifier, torch.cat([noise, fakes]))
        if keep_unfolded:  # parity tests: the oracle (HF ResNetModel) loads the classifier's state_dict as it was BEFORE the fold
            self.classifier_unfolded_state = {k: v.detach().cpu().clone() for k, v in self.classifier.state_dict().items()}
        if self.fold_bn:  # f is frozen and in eval mode: conv + BN is one convolution
            _models.fold_batchnorm(self.classifier)
            if self.channels_last:
                self.classifier.to(memory_format=torch.channels_last)
        if self.device.type == "cuda" and os.environ.get("VCN_FUSE_BIAS_RELU", "0") == "1":
            # frozen f (after the BN fold) and VGG-16: conv bias + ReLU in one vcn_bias_act pass instead of two torch kernels
            if self.fold_bn:
                _models.fuse_conv_bias_relu(self.classifier)
            _models.fuse_conv_bias_relu(self.lpips)
        # frozen f and VGG-16 on the fused bf16 channels_last path (frozen_nets): one launch per conv + bias + ReLU (+ residual),
        # this package's pool / mask kernels between them, no autocast casts.  Same arithmetic as the autocast path it replaces.
        self.fused_frozen = bool(fused_frozen) and self.amp == torch.bfloat16 and self.fold_bn and self.device.type == "cuda"
        self._fused_f = None
        if self.fused_frozen:
            from . import frozen_nets as _fz
            self._fused_f = _fz.FusedResNet(self.classifier)
            if hasattr(self.lpips, "slices"):
                self.lpips.feature_fn = _fz.FusedVGGFeatures(self.lpips, impl=vgg_impl or os.environ.get("VCN_VGG_IMPL", "cudnn"))
        self.deterministic = deterministic
        # bf16 run: synthesis activations stay bf16 NHWC between layers, forward and backward
        self.viscoin_gan.synthesis.resident_bf16 = (self.amp is not None) if self._resident_override is None else bool(self._resident_override)
        self.synthesis_kwargs = {}  # e.g. {'noise_mode': 'const'} for parity runs; the reference passes none (:193)
        _dp.broadcast_parameters([self.classifier, self.concept_extractor, self.explainer, self.viscoin_gan,
                                  self.generator_gan, self.lpips])
        # viscoin.py:130-134; gradients live in flat buckets (dp.GradBuckets) whether or not there are peers
        gan_params = list(self.viscoin_gan.mapping.parameters())
        main_params = list(itertools.chain(self.concept_extractor.parameters(), self.explainer.parameters()))
        # backward reaches the mapping network before Psi: register in forward order, buckets fill in reverse
        if self.buckets_world() > 1 and self.device.type == "cuda" and int(os.environ.get("VCN_DP_RESERVE_SMS", "0")) > 0:
            # opt-in SM partitioning: keep the persistent GEMM grids off the SMs the all-reduce kernels own while they run
            # (bench.py caps NCCL at the same number of CTAs).  Measured on 2 B200s (profiles/r2p_dp_ab.txt): capping NCCL at
            # 4 / 8 / 16 CTAs costs more (17.7 / 15.8 / 15.5 ms) than the contention it removes (15.4 ms unpartitioned) - off.
            from .. import _lib
            _lib.check(_lib.load().vcn_set_conv_sm_reserve(int(os.environ["VCN_DP_RESERVE_SMS"])), "vcn_set_conv_sm_reserve")
        # with peers, the big weights of Psi (psi_fast: bf16 gradients, deposit-only) are exchanged in bf16 - half the bytes of
        # the all-reduce tail (dp.GradBuckets)
        low = []
        if (self.psi_fast and self.amp_psi == torch.bfloat16 and os.environ.get("VCN_DP_BF16_COMM", "1") == "1"
                and all(hasattr(self.concept_extractor, k) for k in ("conv1", "conv2", "conv3", "conv4", "linear1", "linear2"))):
            ce = self.concept_extractor
            low = [ce.conv1.weight, ce.conv2.weight, ce.conv3.weight, ce.conv4.weight, ce.linear1.weight, ce.linear2.weight]
        self.buckets = _dp.GradBuckets(main_params + gan_params, bucket_bytes=int(os.environ.get("VCN_DP_BUCKET_MB", "32")) << 20,
                                       bf16_comm=low)
        self._psi_low = any(c is not None for c in self.buckets.comm)
        fused = (self.device.type == "cuda") if fused_adam is None else fused_adam
        # capturable: step counters and lr live on the device, so the update can sit inside a CUDA graph (step_graphed)
        cap = fused and self.device.type == "cuda"
        lr = torch.tensor(self.p.learning_rate, device=self.device) if cap else self.p.learning_rate
        self.gan_optimizer = torch.optim.Adam(gan_params, lr=lr, fused=fused, capturable=cap)
        self.optimizer = torch.optim.Adam(main_params, lr=lr, fused=fused, capturable=cap)
        self._graphs = {}
        self.lpips_side_stream = os.environ.get("VCN_LPIPS_SIDE_STREAM", "0") == "1" and self.device.type == "cuda"
        self._lpips_stream = None
        self._z_ahead = {}  # per graph key: the z already drawn for the next iteration (step_graphed's generator pipelining)
        self.prefetch_fake = os.environ.get("VCN_PREFETCH_FAKE", "1") == "1"
        self.graph_launches = 0  # libvcn kernel launches replayed from graphs so far
        self._eager_calls = {}

    @staticmethod
    def buckets_world():
        import torch.distributed as dist
        return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    def oracle_state(self):
        """CPU copies of every state_dict the oracle's ``step_losses`` needs (tests / parity tooling; needs keep_unfolded
        when the BatchNorms were folded)."""
        cpu = lambda m: {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
        st = {"classifier": getattr(self, "classifier_unfolded_state", None) or cpu(self.classifier),
              "concept_extractor": cpu(self.concept_extractor), "explainer": cpu(self.explainer),
              "viscoin_gan": cpu(self.viscoin_gan), "generator_gan": cpu(self.generator_gan)}
        st["lpips_convs"] = [(l.weight.detach().cpu().clone(), l.bias.detach().cpu().clone()) for s in self.lpips.slices
                             for l in s if isinstance(l, torch.nn.Conv2d)]
        st["lpips_lins"] = [p.detach().cpu().clone() for p in self.lpips.lins]
        return st

    def trainable_numel(self):
        return self.buckets.numel()

    def _autocast(self):
        return torch.autocast("cuda", dtype=self.amp, enabled=self.amp is not None)

    def _f(self, images):
        if self._fused_f is not None:
            return self._fused_f(images)
        if self.channels_last:
            images = images.contiguous(memory_format=torch.channels_last)
        with self._autocast():
            logits, hidden = self.classifier(images)
        return logits.float(), hidden

    def forward_losses(self, real_images, labels, i, fake_images=None, z_generator=None):
        p = self.p
        self.concept_extractor.train()
        self.explainer.train(not self.deterministic)
        self.viscoin_gan.train()
        if fake_images is None:
            fake_images = synthetic_samples(self.generator_gan, len(real_images), self.device, z_generator)
        all_images = torch.cat([real_images, fake_images.to(real_images.dtype)])
        target_feats = None
        if self.lpips_side_stream and all_images.is_cuda and getattr(self.lpips, "fused", False):
            # (opt-in; measured neutral on one B200, profiles/r2r_ab.txt: 14.58 vs 14.59 ms - the single-stream graph leaves no
            # idle SMs worth filling)
            # The LPIPS features of the TARGET images depend on nothing but the images: a second stream computes them (16 images
            # through VGG-16, dense convolutions) while this one walks f -> Psi -> Theta -> mapping -> the low-resolution
            # synthesis layers, a chain of short launches that leaves SMs idle between and inside kernels.
            if self._lpips_stream is None:
                self._lpips_stream = torch.cuda.Stream(device=self.device)
            main = torch.cuda.current_stream(self.device)
            self._lpips_stream.wait_stream(main)
            with torch.cuda.stream(self._lpips_stream), torch.no_grad():
                src = all_images
                target_feats = self.lpips.features(src)
                for t in target_feats:
                    t.record_stream(main)
        classes, hidden = self._f(all_images)
        if (self.psi_fast and self.amp_psi == torch.bfloat16 and torch.is_grad_enabled()
                and _psi_fast.supported(self.concept_extractor, hidden[-3:])):
            # explicit bf16 path: one-pass weight shadows, gradients deposited straight into the flat buckets (psi_fast.py)
            encoded_concepts, extra_info = _psi_fast.forward(self.concept_extractor, hidden[-3:], self.buckets)
        else:
            if getattr(self, "_psi_low", False) and torch.is_grad_enabled():
                raise RuntimeError("Psi's bf16 communication buckets need the psi_fast path (set VCN_DP_BF16_COMM=0 to train Psi under autocast)")
            with torch.autocast("cuda", dtype=self.amp_psi or torch.bfloat16, enabled=self.amp_psi is not None):
                if self.amp_psi is None and self.amp is not None:
                    hidden = [h.float() for h in hidden[-3:]]
                encoded_concepts, extra_info = self.concept_extractor(hidden[-3:])
        encoded_concepts, extra_info = encoded_concepts.float(), extra_info.float()
        explainer_classes = self.explainer(encoded_concepts)
        losses = {"acc": F.cross_entropy(classes[: len(labels)], labels)}
        if i > p.cd_fid_iteration:
            losses["cr"] = p.delta * concept_regularization_loss(encoded_concepts)
            losses["of"] = p.alpha * output_fidelity_loss(classes, explainer_classes)
        losses["ortho"] = concept_orthogonality_loss(self.concept_extractor)
        rebuilt_images, gan_latents = self.viscoin_gan(z1=encoded_concepts, z2=extra_info, return_latents=True,
                                                       **self.synthesis_kwargs)
        rebuilt_classes, _ = self._f(rebuilt_images)
        if target_feats is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._lpips_stream)
        with self._autocast():
            # the fused feature stack converts layout and dtype itself (one pass, image_ops.image_prep)
            pre_cl = self.channels_last and getattr(self.lpips, "feature_fn", None) is None
            lp = lambda a, b: self.lpips(a.contiguous(memory_format=torch.channels_last) if pre_cl else a,
                                         b.contiguous(memory_format=torch.channels_last) if pre_cl else b,
                                         **({"target_features": target_feats} if target_feats is not None else {}))
            losses["rec"] = reconstruction_loss(rebuilt_images, all_images, rebuilt_classes, classes, lp, p.gamma, p.beta)
        losses["gan"] = gan_regularization_loss(gan_latents, self.viscoin_gan)
        return losses

    def step_graphed(self, real_images, labels, i=1000, z_generator=None, warm=2):
        """step() replayed from ONE CUDA graph: the ~2500 kernel launches of an iteration (library networks, the sm_100a
        synthesis kernels with their tensor maps as kernel parameters, both Adam updates, the bucket memsets and - under
        torch.distributed - the NCCL all-reduces) are captured once per (batch size, i > cd_fid_iteration) and then cost
        one host call.  The first `warm` calls run eagerly (they fill the weight-pack caches and the allocator); z is
        drawn on the host every call like utils.py:37 and copied into the graph's static input.

        Software pipelining of the frozen generator (``self.prefetch_fake``): `synthetic_samples` depends on nothing the
        step updates (generator_gan is frozen, viscoin.py:116,163-166), so the graph generates the fake images of the NEXT
        iteration right after backward - under the tail of the gradient all-reduce, whose last buckets (Psi's convolutions)
        only become final with the last backward kernels - and this iteration consumes the batch the previous replay left
        in a static buffer.  The z stream is unchanged: iteration k still sees the k-th draw of `z_generator`; the draw for
        iteration k+1 just happens one call earlier."""
        n = len(real_images)
        key = (n, i > self.p.cd_fid_iteration)
        z = self._z_ahead.pop(key, None)
        if z is None:
            z = torch.randn(n, self.generator_gan.z_dim, generator=z_generator)
        if self._eager_calls.get(key, 0) < warm and key not in self._graphs:
            # eager warm-up on a SIDE stream (per graph key: every branch a capture will record has run eagerly first):
            # autograd binds each parameter's gradient accumulator to the stream it first runs on, and the legacy default
            # stream cannot take part in a capture
            self._eager_calls[key] = self._eager_calls.get(key, 0) + 1
            if not hasattr(self, "_side"):
                self._side = torch.cuda.Stream(device=self.device)
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                with torch.no_grad():
                    fake = self.generator_gan(z.to(self.device), None)
                out = self.step(real_images.to(self.device, copy=True), labels.to(self.device, copy=True), i, fake_images=fake)
            cur.wait_stream(self._side)
            return out
        if key not in self._graphs:
            from .. import _lib
            st = {"real": torch.empty(real_images.shape, dtype=real_images.dtype, device=self.device),
                  "labels": torch.empty(labels.shape, dtype=labels.dtype, device=self.device),
                  "z": torch.empty(n, self.generator_gan.z_dim, device=self.device)}
            st["real"].copy_(real_images); st["labels"].copy_(labels); st["z"].copy_(z)
            st["prefetch"] = bool(self.prefetch_fake)
            if st["prefetch"]:  # this iteration's batch, generated eagerly once; every replay leaves the next one here
                with torch.no_grad():
                    st["fake"] = self.generator_gan(st["z"], None).clone()
            torch.cuda.synchronize(self.device)
            count0 = _lib.load().vcn_launch_count()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                if st["prefetch"]:
                    def next_fakes():
                        with torch.no_grad():
                            st["fake"].copy_(self.generator_gan(st["z"], None))
                    st["out"] = self.step(st["real"], st["labels"], i, fake_images=st["fake"], _after_backward=next_fakes)
                else:
                    with torch.no_grad():
                        fake = self.generator_gan(st["z"], None)
                    st["out"] = self.step(st["real"], st["labels"], i, fake_images=fake)
            st["graph"] = graph
            st["launches"] = int(_lib.load().vcn_launch_count() - count0)  # libvcn kernels per replay
            self._graphs[key] = st
            if st["prefetch"]:
                # the capture does not execute: st["fake"] still holds G(z) of THIS iteration; fall through to the replay,
                # which consumes it and generates the next one from the z drawn below
                pass
        st = self._graphs[key]
        st["real"].copy_(real_images, non_blocking=True)
        st["labels"].copy_(labels, non_blocking=True)
        if st["prefetch"]:
            z = torch.randn(n, self.generator_gan.z_dim, generator=z_generator)  # iteration k+1's draw
            self._z_ahead[key] = z
        st["z"].copy_(z, non_blocking=True)
        st["graph"].replay()
        self.graph_launches += st["launches"]
        return st["out"]

    def step(self, real_images, labels, i=1000, fake_images=None, z_generator=None, _after_backward=None):
        """One iteration (gradient_accumulation = 1).  Returns the dict of detached loss terms + 'total'.
        `_after_backward` (step_graphed): work that is independent of this step's update, issued between the last
        backward kernel and the wait on the gradient exchange so that it runs under the all-reduce tail."""
        losses = self.forward_losses(real_images, labels, i, fake_images, z_generator)
        total = sum(losses.values())
        total.backward()
        if _after_backward is not None:
            _after_backward()
        self.buckets.finish()
        self.optimizer.step()
        self.gan_optimizer.step()
        self.buckets.reset()
        out = {k: v.detach() for k, v in losses.items()}
        out["total"] = total.detach()
        return out
