"""GPU test of the allocation-free entry point `vcn_bias_act_static` (include/vcn_b200.h): same call as
`vcn_bias_act`, with the reduction scratch of the fused bias gradient in a file-scope `static __device__` array of the
library instead of a caller workspace.  No counterpart in the reference.

Collected last on purpose (`test_zz_*`).

Checked through the C ABI (ctypes), twice in a row (the array keeps the previous call's contents):
  * y and db of the static entry point == y and db of `vcn_bias_act` with a torch-allocated workspace, bit for bit
    (both run the same kernels in the same fixed summation order), and within the bias_act gate (rel-L-inf <= 1e-5 in
    fp32) of the oracle;
  * a shape whose scratch exceeds the built-in array is rejected with VCN_ERR_INVALID and a message;
  * no torch allocation happens during the calls (torch's allocator statistics do not move).
"""

import pytest
import torch

from oracle import ref_ops as R

pytestmark = pytest.mark.gpu

LRELU, F32, BF16 = 3, 0, 2


def _lib():
    from viscoin_next_b200 import _lib
    return _lib


def _call_pair(lib, L, x, b, dy_like, grad, act_id, alpha, gain, clamp, dtype_code):
    """(y, db) from vcn_bias_act with a workspace and from vcn_bias_act_static, same inputs."""
    n, c, h, w = x.shape
    step_b = h * w
    st = L.current_stream(x.device)
    outs = []
    ws_bytes = lib.vcn_bias_act_workspace(dtype_code, x.numel(), c, step_b)
    assert 0 < ws_bytes <= lib.vcn_bias_act_static_scratch_bytes()
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    y0, db0 = torch.empty_like(x), torch.empty(c, dtype=x.dtype, device=x.device)
    y1, db1 = torch.full_like(x, float("nan")), torch.full((c,), float("nan"), dtype=x.dtype, device=x.device)
    yref = dy_like
    rc = lib.vcn_bias_act(L.ptr(x), L.ptr(b), None, L.ptr(yref), None, L.ptr(y0), L.ptr(db0), L.ptr(ws), ws_bytes,
                          dtype_code, grad, act_id, alpha, gain, clamp, x.numel(), c, step_b, st)
    L.check(rc, "vcn_bias_act")
    before = torch.cuda.memory_stats(x.device)["allocation.all.allocated"]
    rc = lib.vcn_bias_act_static(L.ptr(x), L.ptr(b), None, L.ptr(yref), None, L.ptr(y1), L.ptr(db1),
                                 dtype_code, grad, act_id, alpha, gain, clamp, x.numel(), c, step_b, st)
    L.check(rc, "vcn_bias_act_static")
    assert torch.cuda.memory_stats(x.device)["allocation.all.allocated"] == before
    torch.cuda.synchronize()
    return (y0, db0), (y1, db1)


def test_static_scratch_matches_workspace_entry_point_twice():
    L = _lib()
    lib = L.load()
    dev = torch.device("cuda:0")
    L.require_device(0)
    assert lib.vcn_bias_act_static_scratch_bytes() == 1 << 20
    g = torch.Generator().manual_seed(1234)
    # two calls in a row with different shapes and data: the second must not see the first one's partial sums
    for (n, c, h, w), dt, code in (((4, 24, 64, 64), torch.float32, F32), ((3, 16, 96, 48), torch.float32, F32),
                                   ((4, 32, 64, 64), torch.bfloat16, BF16)):
        x = torch.randn(n, c, h, w, generator=g)
        b = torch.randn(c, generator=g)
        xd, bd = x.to(dev, dt), b.to(dev, dt)
        # forward with bias gradient slot filled (db = per-channel sum of y)
        (y0, db0), (y1, db1) = _call_pair(lib, L, xd, bd, None, 0, LRELU, 0.2, 2 ** 0.5, -1.0, code)
        assert torch.equal(y0, y1) and torch.equal(db0, db1)
        y_ref = R.bias_act(xd.float().cpu(), bd.float().cpu(), act="lrelu")
        tol = 1e-5 if dt == torch.float32 else 1e-2
        assert ((y1.float().cpu() - y_ref).abs().max() / y_ref.abs().max().clamp_min(1)).item() <= tol
        s_ref = y1.double().sum(dim=(0, 2, 3)).cpu()
        s_tol = 1e-5 if dt == torch.float32 else 1e-2
        assert ((db1.double().cpu() - s_ref).abs().max() / s_ref.abs().max().clamp_min(1)).item() <= s_tol
        # grad 1: x holds dy, yref the forward output; db = the bias gradient of bias_act.py:173
        dy = torch.randn(n, c, h, w, generator=g).to(dev, dt)
        (dx0, dbg0), (dx1, dbg1) = _call_pair(lib, L, dy, None, y1, 1, LRELU, 0.2, 2 ** 0.5, -1.0, code)
        assert torch.equal(dx0, dx1) and torch.equal(dbg0, dbg1)
        if dt == torch.float32:
            xr = x.clone().requires_grad_(True)
            br = b.clone().requires_grad_(True)
            R.bias_act(xr, br, act="lrelu").backward(dy.float().cpu())
            assert ((dx1.cpu() - xr.grad).abs().max() / xr.grad.abs().max().clamp_min(1)).item() <= 1e-5
            assert ((dbg1.cpu() - br.grad).abs().max() / br.grad.abs().max().clamp_min(1)).item() <= 1e-5


def test_static_scratch_rejects_oversized_shape():
    L = _lib()
    lib = L.load()
    L.require_device(0)
    # [n*c = 2^18+1 rows of 1024] needs (2^18+1)*4 bytes > 1 MiB of partial sums: rejected before any launch, so the
    # pointers are never dereferenced (a small real tensor stands in for them)
    rows, step = (1 << 18) + 1, 1024
    assert lib.vcn_bias_act_workspace(F32, rows * step, rows, step) > lib.vcn_bias_act_static_scratch_bytes()
    t = torch.zeros(16, device="cuda:0")
    rc = lib.vcn_bias_act_static(L.ptr(t), None, None, None, None, L.ptr(t), L.ptr(t), F32, 0, LRELU, 0.2, 1.0, -1.0,
                                 rows * step, rows, step, L.current_stream(t.device))
    assert rc == 1  # VCN_ERR_INVALID
    assert b"built-in array" in lib.vcn_last_error()
    torch.cuda.synchronize()
