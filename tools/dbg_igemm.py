"""Debug helper: vcn_conv_igemm vs torch CPU conv on bf16-rounded operands (isolates the GEMM machinery)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import viscoin_next_b200 as V
from viscoin_next_b200 import _lib
from viscoin_next_b200.host import modconv as M

def run(n, cin, cout, h, w, k, kind):
    g = torch.Generator().manual_seed(n * 1000 + cin + cout + h)
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    wgt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).bfloat16().float()
    if kind == _lib.CONV_FPROP:
        ref = F.conv2d(x, wgt, padding=k // 2)
    elif kind == _lib.CONV_TCONV2:
        ref = F.conv_transpose2d(x, wgt.transpose(0, 1), stride=2)
    else:
        ref = F.conv2d(x, wgt.flip([2, 3]), padding=k // 2)
    xh = M.modulate_pack(x.cuda(), None)
    wt, _ = M.pack_weights(wgt.cuda())
    try:
        y = M.conv_igemm(xh, wt, kind, cout, k, k)
        torch.cuda.synchronize()
    except Exception as e:
        print("FAIL", (n, cin, cout, h, w, k, kind), str(e)[:300], flush=True)
        sys.exit(3)
    err = (y.cpu() - ref).abs().max().item() / ref.abs().max().item()
    print("%-40s rel err %.3e %s" % ((n, cin, cout, h, w, k, kind), err, "ok" if err < 1e-4 else "MISMATCH"), flush=True)
    return err

cases = [(1, 64, 64, 16, 16, 1, 0), (1, 64, 64, 16, 16, 3, 0), (2, 128, 128, 8, 8, 3, 0), (3, 64, 32, 4, 4, 3, 0),
         (1, 8, 6, 8, 8, 3, 0), (2, 72, 40, 9, 9, 3, 0), (1, 128, 256, 32, 32, 3, 0), (2, 512, 512, 16, 16, 3, 0),
         (1, 64, 3, 32, 32, 1, 0), (1, 64, 64, 16, 16, 3, 2), (1, 64, 64, 8, 8, 3, 1), (2, 128, 96, 16, 16, 3, 1),
         (1, 256, 128, 64, 64, 3, 1), (1, 128, 128, 256, 256, 3, 0)]
for c in cases:
    run(*c)
