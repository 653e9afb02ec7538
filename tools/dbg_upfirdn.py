"""Debug helper: run upfirdn2d cases one by one with a sync after each; stop at the first CUDA failure."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import viscoin_next_b200 as V
from oracle import ref_ops as R

f4 = R.setup_filter([1, 3, 3, 1])
CASES = []
for shape in [(2, 3, 8, 8), (2, 3, 9, 9), (1, 5, 17, 17), (2, 4, 32, 32), (1, 3, 64, 64), (1, 2, 128, 128), (2, 3, 16, 16), (1, 2, 65, 65)]:
    for name, kw in dict(A=dict(padding=[1, 1, 1, 1], gain=4), Ab=dict(padding=[2, 2, 2, 2], flip_filter=True, gain=4),
                         B=dict(up=2, padding=[2, 1, 2, 1], gain=4), Bb=dict(down=2, padding=[1, 1, 1, 1], flip_filter=True, gain=4)).items():
        CASES.append((shape, name, kw))
only = int(sys.argv[1]) if len(sys.argv) > 1 else None
for i, (shape, name, kw) in enumerate(CASES):
    if only is not None and i != only:
        continue
    x = torch.randn(shape, generator=torch.Generator().manual_seed(i))
    ref = R.upfirdn2d(x, f4, **kw)
    try:
        y = V.upfirdn2d.upfirdn2d(x.cuda(), f4.cuda(), **kw)
        torch.cuda.synchronize()
        err = (y.cpu() - ref).abs().max().item()
        print(i, shape, name, "ok" if err < 1e-5 else "MISMATCH", err, flush=True)
    except Exception as e:
        print(i, shape, name, "CUDA FAIL", str(e)[:200], flush=True)
        open("gpurun_out/dbg_first_fail.txt", "w").write(str(i))
        sys.exit(3)
