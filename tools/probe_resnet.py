import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import viscoin_next_b200 as V
import viscoin_next_b200.host.frozen_nets as fz
VM = V.viscoin_models
torch.manual_seed(0)
cls = VM.Classifier(200).cuda().eval().requires_grad_(False)
VM.calibrate_batchnorm(cls, torch.randn(8, 3, 256, 256).cuda())
VM.fold_batchnorm(cls)
def rel(a, b): return float((a.float() - b.float()).abs().max() / b.float().abs().max())
x = torch.randn(4, 3, 256, 256, device="cuda")
r = cls.resnet
# embedder
e_ref = r.embedder(x)
xb = x.bfloat16().contiguous(memory_format=torch.channels_last)
e = F.max_pool2d(fz.conv_relu(xb, r.embedder.embedder.convolution), 3, 2, 1)
print("embedder", rel(e, e_ref))
h_ref, h = e_ref, e
for si, stage in enumerate(r.encoder.stages):
    for li, layer in enumerate(stage.layers):
        # per-op check on the SAME input (h_ref) to isolate op errors
        inp = h_ref.bfloat16().contiguous(memory_format=torch.channels_last)
        convs = list(layer.layer)
        a_ref = convs[0](h_ref); a = fz.conv_relu(inp, convs[0].convolution)
        b_ref = convs[1](a_ref); b = fz.conv_relu(a_ref.bfloat16().contiguous(memory_format=torch.channels_last), convs[1].convolution)
        c_ref = convs[2](b_ref)
        sc = layer.shortcut
        res_ref = sc(h_ref)
        res = inp if isinstance(sc, torch.nn.Identity) else fz.conv_relu(inp, sc.convolution, relu=False)
        out_ref = torch.relu(c_ref + res_ref)
        out = fz.conv_relu(b_ref.bfloat16().contiguous(memory_format=torch.channels_last), convs[2].convolution, relu=True, residual=res_ref.bfloat16().contiguous(memory_format=torch.channels_last))
        full = fz.FusedResNet(cls)._bottleneck(layer, h)
        full_ref = layer(h_ref)
        print("stage %d layer %d: conv1 %.1e conv2 %.1e shortcut %.1e add_relu %.1e | chained %.1e  (|c| %.2f |res| %.2f |out| %.2f)" % (
            si, li, rel(a, a_ref), rel(b, b_ref), rel(res, res_ref), rel(out, out_ref), rel(full, full_ref),
            float(c_ref.abs().max()), float(res_ref.abs().max()), float(out_ref.abs().max())))
        h_ref, h = full_ref, full
