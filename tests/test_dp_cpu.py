"""Data-parallel gradient exchange (viscoin_next_b200.dp.GradBuckets) on CPU: world_size 2 over gloo.

The reference has no test for its distributed path; the property checked here is the definition of data
parallelism: after the exchange every rank holds the gradient of the mean loss over the union of the shards, i.e. the
same numbers a single process computes on the concatenated batch."""

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _model():
    torch.manual_seed(7)
    return torch.nn.Sequential(torch.nn.Linear(12, 33), torch.nn.ReLU(), torch.nn.Linear(33, 5), torch.nn.ReLU(),
                               torch.nn.Linear(5, 3))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from viscoin_next_b200 import dp
    m = _model()
    if rank != 0:  # replicas start different; broadcast_parameters must make them identical
        with torch.no_grad():
            for p in m.parameters():
                p.add_(1.0)
    versions = [p._version for p in m.parameters()]
    dp.broadcast_parameters([m])
    # the in-place broadcast must bump every tensor's version counter: the synthesis path's packed-weight / style / noise
    # caches are keyed on (data_ptr, _version) and may have been filled before the broadcast (dp.py)
    assert rank == 0 or all(p._version > v for p, v in zip(m.parameters(), versions)), "broadcast did not bump tensor versions"
    buckets = dp.GradBuckets(m.parameters(), bucket_bytes=512)  # several small buckets
    assert len(buckets.buckets) >= 3
    g = torch.Generator().manual_seed(100)
    x = torch.randn(8, 12, generator=g)
    xs = x[rank * 4:(rank + 1) * 4]
    res = []
    for step in range(2):
        m(xs).square().mean().backward()
        buckets.finish()
        res.append([p.grad.clone() for p in m.parameters()])
        buckets.reset()
        assert all(float(p.grad.abs().sum()) == 0 for p in m.parameters())
    torch.save(res, os.path.join(out_dir, "r%d.pt" % rank))
    dist.destroy_process_group()


def test_bucketed_allreduce_matches_single_process(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(os.path.join(tmp_path, "r0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "r1.pt"))
    m = _model()
    x = torch.randn(8, 12, generator=torch.Generator().manual_seed(100))
    m(x).square().mean().backward()
    want = [p.grad for p in m.parameters()]
    for step in range(2):
        for a, b, w in zip(r0[step], r1[step], want):
            assert torch.equal(a, b)
            assert torch.allclose(a, w, rtol=1e-5, atol=1e-7)


def test_single_process_buckets_are_views():
    from viscoin_next_b200 import dp
    m = _model()
    buckets = dp.GradBuckets(m.parameters(), bucket_bytes=1 << 20)
    assert len(buckets.buckets) == 1
    m(torch.randn(4, 12)).sum().backward()
    buckets.finish()
    flat = buckets.buckets[0]
    for p in m.parameters():
        assert p.grad.data_ptr() >= flat.data_ptr() and p.grad.data_ptr() < flat.data_ptr() + flat.numel() * 4
    assert float(flat.abs().sum()) > 0
    buckets.reset()
    assert float(flat.abs().sum()) == 0


def _worker_bf16(rank, world, port, out_dir):
    """bf16 communication buckets: the first Linear's weight goes through psi_fast._ShadowWeight (bf16 shadow forward,
    bf16 gradient deposited), is exchanged in bf16 and widened into its fp32 .grad view by finish()."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from viscoin_next_b200 import dp
    from viscoin_next_b200.host import psi_fast
    m = _model()
    w0 = m[0].weight
    buckets = dp.GradBuckets(m.parameters(), bucket_bytes=512, bf16_comm=[w0])
    assert sum(c is not None for c in buckets.comm) == 1 and w0 in buckets._comm_view
    x = torch.randn(8, 12, generator=torch.Generator().manual_seed(100))
    xs = x[rank * 4:(rank + 1) * 4]
    res = []
    for step in range(2):
        h = torch.nn.functional.linear(xs.bfloat16(), psi_fast._ShadowWeight.apply(w0, buckets), m[0].bias.bfloat16()).float()
        m[2:](torch.relu(h)).square().mean().backward()
        buckets.finish()
        res.append([p.grad.clone() for p in m.parameters()])
        buckets.reset()
    torch.save(res, os.path.join(out_dir, "b%d.pt" % rank))
    dist.destroy_process_group()


def test_bf16_communication_bucket(tmp_path):
    port = _free_port()
    mp.spawn(_worker_bf16, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(os.path.join(tmp_path, "b0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "b1.pt"))
    # single process, same arithmetic (bf16 first layer), concatenated batch
    m = _model()
    x = torch.randn(8, 12, generator=torch.Generator().manual_seed(100))
    w16 = m[0].weight.detach().bfloat16().requires_grad_(True)
    h = torch.nn.functional.linear(x.bfloat16(), w16, m[0].bias.bfloat16()).float()
    m[2:](torch.relu(h)).square().mean().backward()
    want = [w16.grad.float()] + [p.grad for p in list(m.parameters())[1:]]
    for step in range(2):
        for i, (a, b, w) in enumerate(zip(r0[step], r1[step], want)):
            assert torch.equal(a, b)
            # the weight gradient is a bf16 mean of two bf16 shard gradients: a few bf16 ulps of the largest entry
            tol = 2e-2 * float(w.abs().max()) if i <= 1 else 1e-5 + 1e-3 * float(w.abs().max())
            assert float((a - w).abs().max()) <= tol, (step, i, float((a - w).abs().max()), tol)
