"""Oracle: functional CPU restatement of one VisCoIN training iteration.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Everything is driven by plain ``state_dict``s with the
reference's key names.  Citations are relative to ``/root/reference``.

  concept_extractor   viscoin/models/concept_extractors.py:102-141
  explainer           viscoin/models/explainers.py:42-56
  losses              viscoin/training/losses.py:25-34,66-96,125-157,165-192
  lpips_vgg           third-party ``lpips`` v0.1.4 (flake.nix:33-42; NOT installable here - restated from the published
                      algorithm: ScalingLayer -> VGG-16 relu1_2..relu5_3 -> normalize_tensor (eps 1e-10) -> squared
                      difference -> 1x1 'lin' -> spatial mean -> sum).  Parity of this term against the real package is
                      UNPINNED; every other term is pinned against the reference modules by make_golden.py.
  classifier          third-party ``transformers.ResNetModel(ResNetConfig())`` + Linear (classifiers.py:83-92); the HF
                      module is the same library code on both sides (like ATen's convolution)
  step_losses         viscoin/training/viscoin.py:163-209 (i > cd_fid_iteration: all six terms)
"""

import torch
import torch.nn.functional as F

from . import ref_net as RN

VGG_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512]
VGG_TAPS = [2, 4, 7, 10, 13]  # number of convolutions before relu1_2, relu2_2, relu3_3, relu4_3, relu5_3


def concept_extractor(sd, latent_layers, p=""):
    assert len(latent_layers) == 3
    relu = torch.relu
    x1 = relu(F.conv2d(latent_layers[0], sd[p + "conv1.weight"], sd[p + "conv1.bias"], stride=4, padding=2))
    x2 = relu(F.conv2d(latent_layers[1], sd[p + "conv2.weight"], sd[p + "conv2.bias"], stride=2, padding=1))
    x3 = relu(F.conv2d(latent_layers[2], sd[p + "conv3.weight"], sd[p + "conv3.bias"], stride=1, padding=1))
    x = relu(F.conv2d(torch.cat((x1, x2, x3), 1), sd[p + "conv4.weight"], sd[p + "conv4.bias"], padding=1))
    y1 = F.adaptive_avg_pool2d(relu(F.conv2d(x, sd[p + "conv5.weight"], sd[p + "conv5.bias"])), 3)
    y2 = F.adaptive_avg_pool2d(x, 3).reshape(x.shape[0], -1)
    y2 = relu(F.linear(y2, sd[p + "linear1.weight"], sd[p + "linear1.bias"]))
    y2 = relu(F.linear(y2, sd[p + "linear2.weight"]))
    return y1, y2


def explainer(sd, x, p=""):
    """Dropout(0.01) is the identity here: parity runs keep Theta's dropout off (RNG streams differ per device)."""
    return F.linear(F.adaptive_max_pool2d(x, 1).flatten(start_dim=1), sd[p + "linear.weight"], sd[p + "linear.bias"])


def cross_cross_entropy_loss(prediction, target_prediction):
    p = F.softmax(prediction, dim=1)
    t = F.softmax(target_prediction, dim=1)
    return (p.log() * -t).sum(dim=1).mean()


def l1_loss(x):
    return F.l1_loss(x, torch.zeros(x.shape))


def concept_regularization_loss(concept_embeddings):
    pooled = F.adaptive_max_pool2d(concept_embeddings, 1).flatten(start_dim=1)
    return l1_loss(F.normalize(pooled, p=2, dim=1)) + l1_loss(concept_embeddings)


def concept_orthogonality_loss(conv5_weight):
    k = conv5_weight.shape[0]
    nw = F.normalize(conv5_weight.reshape(k, -1), dim=1).abs()
    return ((nw @ nw.T).sum() - k) / (k ** 2)


def vgg_features(convs, x):
    """convs: the 13 (weight, bias) pairs of VGG-16 in order.  Returns the 5 LPIPS taps."""
    out, ci = [], 0
    for v in VGG_CFG:
        if v == "M":
            x = F.max_pool2d(x, 2, 2)
            continue
        w, b = convs[ci]
        x = torch.relu(F.conv2d(x, w, b, padding=1))
        ci += 1
        if ci in VGG_TAPS:
            out.append(x)
    return out


def lpips_vgg(convs, lins, in0, in1):
    shift = torch.tensor([-0.030, -0.088, -0.188]).view(1, 3, 1, 1)
    scale = torch.tensor([0.458, 0.448, 0.450]).view(1, 3, 1, 1)
    f0 = vgg_features(convs, (in0 - shift) / scale)
    f1 = vgg_features(convs, (in1 - shift) / scale)

    def unit(f, eps=1e-10):
        return f / (torch.sqrt(torch.sum(f ** 2, dim=1, keepdim=True)) + eps)

    val = 0
    for a, b, lin in zip(f0, f1, lins):
        d = (unit(a) - unit(b)) ** 2
        val = val + F.conv2d(d, lin.reshape(1, -1, 1, 1)).mean([2, 3], keepdim=True)
    return val


def reconstruction_loss(reconstructed, original, reconstructed_classes, original_classes, lpips_fn, lambda_classes=0.1,
                        lambda_lpips=3.0):
    return (F.l1_loss(reconstructed, original) + F.mse_loss(reconstructed, original)
            + lambda_classes * cross_cross_entropy_loss(reconstructed_classes, original_classes.detach())
            + lambda_lpips * torch.mean(lpips_fn(reconstructed, original)))


def make_classifier(sd):
    from transformers import ResNetConfig, ResNetModel
    m = ResNetModel(ResNetConfig())
    m.load_state_dict({k[len("resnet."):]: v for k, v in sd.items() if k.startswith("resnet.")})
    m.eval().requires_grad_(False)

    def f(x):
        r = m(x, output_hidden_states=True)
        return F.linear(r["pooler_output"].reshape(x.shape[0], -1), sd["linear.weight"], sd["linear.bias"]), r["hidden_states"]
    return f


def step_losses(state, real_images, labels, fake_images, noise_mode="const", alpha=0.5, beta=3.0, gamma=0.1, delta=0.2):
    """viscoin.py:163-209 with all six terms.  state: dict of state_dicts {'classifier','concept_extractor','explainer',
    'viscoin_gan','lpips_convs','lpips_lins'}; tensors of 'concept_extractor' / 'explainer' / 'viscoin_gan' mapping that
    should receive gradients must have requires_grad set by the caller.  Returns the dict of loss terms."""
    f = state.get("_f") or make_classifier(state["classifier"])
    all_images = torch.cat([real_images, fake_images])
    classes, hidden = f(all_images)
    phi, phi2 = concept_extractor(state["concept_extractor"], hidden[-3:])
    expl = explainer(state["explainer"], phi)
    out = {"acc": F.cross_entropy(classes[: len(labels)], labels),
           "cr": delta * concept_regularization_loss(phi),
           "of": alpha * cross_cross_entropy_loss(expl, classes.detach()),
           "ortho": concept_orthogonality_loss(state["concept_extractor"]["conv5.weight"])}
    rebuilt, ws = RN.generator_adapted(phi, phi2, state["viscoin_gan"], training=True, return_latents=True,
                                       noise_mode=noise_mode)
    rebuilt_classes, _ = f(rebuilt)
    lp = lambda a, b: lpips_vgg(state["lpips_convs"], state["lpips_lins"], a, b)
    out["rec"] = reconstruction_loss(rebuilt, all_images, rebuilt_classes, classes, lp, gamma, beta)
    w_avg = state["viscoin_gan"]["mapping.fixed_w_avg"].detach()
    out["gan"] = F.mse_loss(ws, w_avg.repeat([ws.shape[0], ws.shape[1], 1]))
    out["_rebuilt"] = rebuilt
    return out
