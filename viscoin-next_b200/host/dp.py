"""dp - data-parallel gradient exchange for the VisCoIN train step: one process per GPU, one bucketed
all-reduce (mean) per step over NCCL, launched from autograd hooks so that it overlaps the rest of backward.

The reference trains on a single device (viscoin/training/viscoin.py:84-223 has no collective at all); the only
distributed code in the reference tree is StyleGAN2-ADA's per-parameter ``all_reduce`` after backward
(stylegan2_ada/training/training_loop.py:300-307, not overlapped).  Here:

  * the gradients of all trainable parameters live in a few flat fp32 buffers ("buckets", filled in reverse
    registration order = roughly the order backward produces them); ``p.grad`` is a view into its bucket, so autograd
    accumulates straight into the communication buffer and no gather / scatter copies exist;
  * a post-accumulate hook counts the parameters of a bucket that are ready; the last one launches
    ``all_reduce(bucket, async_op=True)`` (NCCL runs it on its own stream, concurrently with the remaining backward
    kernels); ``finish()`` waits for the handles before the optimizer reads the gradients.

Works with any torch.distributed backend (tests use gloo with world_size 2 on CPU).
"""

import torch
import torch.distributed as dist


class GradBuckets:
    """`bf16_comm`: parameters whose gradient only ever arrives through ``deposit`` as a bf16 tensor (psi_fast: the big
    weights of Psi, whose gradients are bf16 values to begin with).  With peers they get buckets of their own that are
    EXCHANGED in bf16 - half the bytes of the step's all-reduce, whose tail (these very gradients become final with the last
    backward kernels) is the exposed part of the collective - and are widened into the fp32 `p.grad` views by ``finish``.
    Tolerance: one bf16 rounding of the mean per element and NCCL's bf16 summation (tools/dp_check.py gates the exchange at
    rel-L2 <= 5e-3 against the fp32 mean of the shard gradients)."""

    def __init__(self, params, bucket_bytes=64 << 20, group=None, bf16_comm=()):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.buckets = []      # flat fp32 tensors
        self.comm = []         # per bucket: flat bf16 communication buffer or None
        self.members = []      # list of lists of params
        self._pending = []     # per bucket: params still to arrive in this backward
        self._handles = []
        self._bucket_of = {}
        self._comm_view = {}   # param -> bf16 view of its slice of the communication buffer
        low = set(bf16_comm) if self.world > 1 else set()
        cur, cur_bytes, cur_low = [], 0, False
        for p in reversed(self.params):
            assert p.dtype == torch.float32, "trainable parameters are fp32 master weights"
            nbytes = p.numel() * 4
            is_low = p in low
            if cur and (cur_bytes + nbytes > bucket_bytes or is_low != cur_low):
                self._seal(cur, cur_low)
                cur, cur_bytes = [], 0
            cur.append(p)
            cur_bytes += nbytes
            cur_low = is_low
        if cur:
            self._seal(cur, cur_low)
        for p in self.params:
            p.register_post_accumulate_grad_hook(self._on_grad)
        self.reset()

    def _seal(self, plist, low=False):
        dev = plist[0].device
        total = sum((p.numel() + 63) // 64 * 64 for p in plist)  # 256-byte aligned slices
        flat = torch.zeros(total, dtype=torch.float32, device=dev)
        comm = torch.zeros(total, dtype=torch.bfloat16, device=dev) if low else None
        off = 0
        for p in plist:
            p.grad = flat[off:off + p.numel()].view_as(p)
            if low:
                self._comm_view[p] = comm[off:off + p.numel()].view_as(p)
            self._bucket_of[p] = len(self.buckets)
            off += (p.numel() + 63) // 64 * 64
        self.buckets.append(flat)
        self.comm.append(comm)
        self.members.append(list(plist))

    def reset(self):
        """Zero the gradients (replaces optimizer.zero_grad(): the views must stay attached to the buckets)."""
        for flat, comm in zip(self.buckets, self.comm):
            if comm is not None:
                comm.zero_()  # the fp32 views are overwritten by finish(); a gradient that never arrives must read as zero
            else:
                flat.zero_()
        self._pending = [len(m) for m in self.members]
        self._handles = []
        self._fresh = set(self.params)  # parameters whose slice still holds the zeros written by reset()

    def deposit(self, p, g):
        """Hand over the gradient of `p` from outside autograd's accumulation (psi_fast._ShadowWeight): `g` may have another
        dtype / memory format; the first contribution of a step is written into the (zeroed) bucket slice with one fused
        cast-copy, later ones are added.  The caller's backward returns None for `p`: autograd still runs the parameter's
        AccumulateGrad node (a no-op on an undefined gradient) and fires the post-accumulate hook AFTER this call - that hook is
        the parameter's arrival for the bucket's all-reduce, exactly as on the ordinary path."""
        with torch.no_grad():
            dst = self._comm_view.get(p, p.grad)  # bf16 communication slice (peers) or the fp32 gradient view
            if p in self._fresh:
                self._fresh.discard(p)
                dst.copy_(g)
            else:
                dst.add_(g.to(dst.dtype))

    def _on_grad(self, p):
        b = self._bucket_of[p]
        self._pending[b] -= 1
        if self._pending[b] == 0 and self.world > 1:
            self._launch(b)

    def _launch(self, b):
        flat = self.comm[b] if self.comm[b] is not None else self.buckets[b]
        backend = dist.get_backend(self.group)
        if backend == "nccl":
            self._handles.append((dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True), None))
        else:
            self._handles.append((dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True), flat))

    def finish(self):
        """Wait for the exchange of this step; parameters whose gradient never arrived (unused in this step) are
        exchanged now so that every rank issues the same collectives."""
        if self.world > 1:
            for b, left in enumerate(self._pending):
                if left > 0:
                    self._launch(b)
                    self._pending[b] = 0
            for h, flat in self._handles:
                h.wait()
                if flat is not None:
                    flat.div_(self.world)
            for flat, comm in zip(self.buckets, self.comm):
                if comm is not None:
                    flat.copy_(comm)  # widen the exchanged bf16 means into the fp32 views the optimizer reads
        self._handles = []

    def numel(self):
        return sum(p.numel() for p in self.params)


def broadcast_parameters(modules, src=0, group=None):
    """Identical initial replicas: rank `src`'s parameters and buffers to every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    # Receivers copy the broadcast value in with `copy_` on the tensor itself: that bumps `t._version` (a c10d collective
    # writing through `.data` - or into the tensor directly - does not), and the packed-weight, batched-affine and const-noise
    # caches of the synthesis path are keyed on (data_ptr, _version) and may already have been filled by a forward that ran
    # before the broadcast (BatchNorm calibration in VisCoINTrainer.__init__).
    rank = dist.get_rank(group)
    with torch.no_grad():
        for m in modules:
            for t in list(m.parameters()) + list(m.buffers()):
                if rank == src:
                    dist.broadcast(t.detach(), src=src, group=group)
                else:
                    buf = torch.empty_like(t)
                    dist.broadcast(buf, src=src, group=group)
                    t.copy_(buf)
