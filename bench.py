#!/usr/bin/env python
"""bench.py - the measurement contract of this repo (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload trainstep|synthesis|opsweep] [--impl ours|reference|reference-gpu]

One "step" = one pass of the hot path over one batch of synthetic input.  Prints ONE JSON line (rank 0).
  value     whole-job throughput with inputs resident in HBM (device-timed, CUDA events, max over ranks)
  e2e       same metric through the public operator API with HOST (pinned) buffers: H2D + op + D2H timed
  roofline  dominant kernel: algorithmic bytes per launch / CUDA-event duration vs MEASURED_PEAKS.json
  cpu_baseline  the oracle (CPU restatement of the reference ref path) timed on a bounded sample, rank 0, N=1
  roofline_others  every other hand-written kernel alone at its BASELINE shape (+ the cuDNN kernel of the same layer)
`--impl reference` times the reference's CPU implementation (the oracle port; /root/reference does not exist on
the GPU box) on the same workload - the FULL batch, both Adam updates included - with all host threads.
`--impl reference-gpu` (secondary baseline) runs the UNMODIFIED reference modules staged in the
git-ignored oracle/_ref (its own CUDA plugins JIT-built for sm_100a + cuDNN convolutions, fp32/TF32, eager) on the GPU.
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-gpu"])
    ap.add_argument("--workload", default=os.environ.get("VCN_BENCH_WORKLOAD", "auto"))
    ap.add_argument("--batch", type=int, default=8, help="images per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sus=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [s.strip() for s in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx = float(p[2])
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower() == "active":
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # opt-in SM partitioning (VCN_DP_RESERVE_SMS = k): NCCL capped at k CTAs on SMs the persistent GEMM grids leave free
        if int(os.environ.get("VCN_DP_RESERVE_SMS", "0")) > 0:
            os.environ.setdefault("NCCL_MAX_CTAS", os.environ["VCN_DP_RESERVE_SMS"])
        backend = "nccl" if torch.cuda.is_available() and args.impl == "ours" else "gloo"
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, device_id=torch.device("cuda", local) if backend == "nccl" else None)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(ms, world, device):
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return ms



def time_kernel_flushed(fn, device, iters=20):
    """Average device time [ms] of ONE launch of fn() with L2 flushed before every launch.  The launch is replayed from a
    CUDA graph: the python + ctypes + cuTensorMapEncode path costs 30-60 us per call, more than the 256 MB flush takes,
    so an eagerly issued launch would be timed together with the gap the host leaves in front of it."""
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    for _ in range(3):
        fn()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    g.replay()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    return sum(times) / len(times)


# ----------------------------------------------------------------------------------------------------------
# workloads
# ----------------------------------------------------------------------------------------------------------

GATE_SET = [(512, 64), (256, 128), (128, 256)]  # (channels, output resolution) of SURVEY 8(d) config 2: >> L2 at N=8


class OpSweep:
    """BASELINE.json configs[1]: upfirdn2d + bias_act over the memory-bound gate set (fp32, NCHW, N=batch).

    Per step and per (C, H): pattern A FIR ([N,C,H+1,H+1] -> [N,C,H,H], 4x4, pad 1, gain 4: what follows every
    transposed conv of the synthesis network) then bias_act(lrelu) on the result, forward; then
    the bias_act grad-1 (+db) pass on the same tensor.  Algorithmic bytes: upfirdn2d (in+out)*4,
    bias_act fwd 2*S*4, grad-1 3*S*4."""

    name = "upfirdn2d+bias_act op sweep fp32 NCHW N=%d, C x H in {512x64, 256x128, 128x256}, fwd + bias_act grad1"
    metric = "upfirdn2d+bias_act algorithmic HBM GB/s"
    unit = "GB/s"
    dtype = "f32"

    def __init__(self, batch):
        self.n = batch
        self.name = self.name % batch

    def alg_bytes(self, n=None):
        n = self.n if n is None else n
        tot = 0
        for c, h in GATE_SET:
            s_in, s_out = n * c * (h + 1) * (h + 1), n * c * h * h
            tot += (s_in + s_out) * 4 + 2 * s_out * 4 + 3 * s_out * 4
        return tot

    def units_per_step(self):
        return self.alg_bytes() / 1e9

    def setup(self, device, seed):
        import viscoin_next_b200 as V
        self.V = V
        g = torch.Generator(device=device).manual_seed(seed)
        self.f = V.upfirdn2d.setup_filter([1, 3, 3, 1], device=device)
        self.x = [torch.randn(self.n, c, h + 1, h + 1, device=device, generator=g) for c, h in GATE_SET]
        self.b = [torch.randn(c, device=device, generator=g) for c, _ in GATE_SET]
        self.dy = [torch.randn(self.n, c, h, h, device=device, generator=g) for c, h in GATE_SET]

    def _one(self, x, b, dy):
        V = self.V
        y = V.upfirdn2d.upfirdn2d(x, self.f, padding=[1, 1, 1, 1], gain=4)
        y = y.requires_grad_(True)
        b = b.requires_grad_(True)
        z = V.bias_act.bias_act(y, b, act="lrelu")
        dx, db = torch.autograd.grad(z, [y, b], dy)
        return dx, db

    def _sweep(self):
        out = None
        for x, b, dy in zip(self.x, self.b, self.dy):
            out = self._one(x, b, dy)
        return out

    def step(self):
        """Device-timed sweep.  The nine launches of a sweep take ~0.62 ms on the GPU and about as long to issue from python
        (ctypes + autograd): like the other workloads the timed region replays them from ONE CUDA graph (captured after two
        eager sweeps; VCN_BENCH_GRAPH=0 times the eager launches); `e2e` stays eager with host buffers."""
        if os.environ.get("VCN_BENCH_GRAPH", "1") != "1":
            return self._sweep()
        self._calls = getattr(self, "_calls", 0) + 1
        if self._calls <= 2:
            return self._sweep()
        if getattr(self, "_graph", None) is None:
            lib = self.V._lib.load()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._sweep()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            c0 = lib.vcn_launch_count()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._gout = self._sweep()
            self._glaunches = int(lib.vcn_launch_count() - c0)
        self._graph.replay()
        self.graph_launches = getattr(self, "graph_launches", 0) + self._glaunches
        return self._gout

    # --- end to end: host buffers in, host result out -------------------------------------------------
    def e2e_setup(self, device):
        self.hx = [x.cpu().pin_memory() for x in self.x]
        self.hdy = [d.cpu().pin_memory() for d in self.dy]
        self.hout = [torch.empty(d.shape, dtype=d.dtype).pin_memory() for d in self.dy]
        self.h2d = sum(t.numel() * 4 for t in self.hx + self.hdy)
        self.d2h = sum(t.numel() * 4 for t in self.hout)

    def e2e_step(self, device):
        for hx, hdy, b, ho in zip(self.hx, self.hdy, self.b, self.hout):
            dx, _ = self._one(hx.to(device, non_blocking=True), b, hdy.to(device, non_blocking=True))
            ho.copy_(dx, non_blocking=True)

    # --- dominant kernel alone ----------------------------------------------------------------------
    def roofline(self, device, peaks, iters=20):
        """bias_act forward (rows kernel) on the largest tensor [N,128,256,256]: 2*S*4 algorithmic bytes."""
        V = self.V
        c, h = GATE_SET[-1]
        x = torch.randn(self.n, c, h, h, device=device)
        b = self.b[-1].detach()
        ms = time_kernel_flushed(lambda: V.bias_act.bias_act(x, b, act="lrelu"), device, iters)
        alg = 2 * x.numel() * 4
        ach = alg / (ms * 1e-3) / 1e9
        # the FIR kernels of the same gate shape, same method (reported next to the dominant kernel, same peak)
        others = []
        f = self.f
        xa = torch.randn(self.n, c, h + 1, h + 1, device=device)
        for label, xin in (("upfirdn2d_tile_kernel<float> pattern A [%d,%d,%d,%d] fp32 NCHW" % (self.n, c, h + 1, h + 1), xa),
                           ("upfirdn2d_cl_sep_kernel<bf16> pattern A [%d,%d,%d,%d] bf16 NHWC" % (self.n, c, h + 1, h + 1),
                            xa.bfloat16().contiguous(memory_format=torch.channels_last))):
            t = time_kernel_flushed(lambda: V.upfirdn2d.upfirdn2d(xin, f, padding=[1, 1, 1, 1], gain=4), device, iters)
            nbytes = (xin.numel() + self.n * c * h * h) * xin.element_size()
            a_ = nbytes / (t * 1e-3) / 1e9
            others.append({"kernel": label, "bound": "hbm", "achieved": round(a_, 1), "peak": peaks["hbm"], "unit": "GB/s",
                           "frac": round(a_ / peaks["hbm"], 4), "alg_bytes_per_launch": nbytes, "avg_launch_ms": round(t, 5)})
        self.other_rooflines = others
        return {"kernel": "bias_act_rows_kernel<float,lrelu> [%d,%d,%d,%d] fwd" % (self.n, c, h, h), "bound": "hbm",
                "achieved": round(ach, 1), "peak": peaks["hbm"], "unit": "GB/s", "frac": round(ach / peaks["hbm"], 4),
                "peak_source": peaks["src"] + " (burst copy)", "traffic": None, "alg_bytes_per_launch": alg,
                "avg_launch_ms": round(ms, 5)}

    def reference_gpu_setup(self, device, seed):
        """The reference's own plugins (bias_act.cu:23-147, upfirdn2d.cu:97-200) through its operator API, impl='cuda'."""
        from oracle import stage_ref
        import types
        ref = stage_ref.load_plugins()
        self.setup(device, seed)
        self.V = types.SimpleNamespace(upfirdn2d=ref.upfirdn2d, bias_act=ref.bias_act)
        self.f = ref.upfirdn2d.setup_filter([1, 3, 3, 1], device=device)
        return "reference plugins bias_act.cu / upfirdn2d.cu (oracle/_ref, unmodified, sm_100a JIT) through torch_utils.ops, impl='cuda'"

    # --- CPU (oracle) -------------------------------------------------------------------------------
    def cpu_step(self, n_sample, seed=0):
        """The oracle's restatement of the reference ref path on a sample of n_sample images per tensor."""
        from oracle import ref_ops as R
        g = torch.Generator().manual_seed(seed)
        f = R.setup_filter([1, 3, 3, 1])
        t0 = time.perf_counter()
        for c, h in GATE_SET:
            x = torch.randn(n_sample, c, h + 1, h + 1, generator=g)
            b = torch.randn(c, generator=g).requires_grad_(True)
            dy = torch.randn(n_sample, c, h, h, generator=g)
            t0 -= 0  # generation is outside the op but cheap; keep the timer simple and conservative
            y = R.upfirdn2d(x, f, padding=[1, 1, 1, 1], gain=4).requires_grad_(True)
            z = R.bias_act(y, b, act="lrelu")
            torch.autograd.grad(z, [y, b], dy)
        return time.perf_counter() - t0


class Synthesis:
    """BASELINE.json configs[0] on the GPU: StyleGAN2-ADA SynthesisNetwork forward, 256x256, batch N, random-init,
    fp32 storage (bf16 tensor-core operands), noise_mode='const' with noise_strength 0.1.
    One step = ws[N,14,512] -> img[N,3,256,256] through this package's modules (every layer one fused sm_100a call)."""

    metric = "G-synthesis images/s @256^2"
    unit = "images/s"
    dtype = "bf16"
    RES = 256

    def __init__(self, batch):
        self.n = batch
        self.name = "SynthesisNetwork forward %dx%d batch %d/GPU, random-init, noise const 0.1, fp32 I/O, bf16-resident activations + bf16 MMA operands" % (self.RES, self.RES, batch)

    def units_per_step(self):
        return float(self.n)

    def _build(self, device, N):
        torch.manual_seed(0)
        S = N.SynthesisNetwork(w_dim=512, img_resolution=self.RES, img_channels=3)
        with torch.no_grad():
            for name, p in S.named_parameters():
                if name.endswith("noise_strength"):
                    p.fill_(0.1)
        S = S.to(device).eval().requires_grad_(False)
        S.resident_bf16 = device.type == "cuda"  # bf16 NHWC activations between layers (stated in `config.workload`)
        return S

    def setup(self, device, seed):
        import viscoin_next_b200 as V
        self.V = V
        self.S = self._build(device, V.networks)
        g = torch.Generator(device=device).manual_seed(seed)
        self.ws = torch.randn(self.n, self.S.num_ws, 512, device=device, generator=g)
        self.graphed = None
        if os.environ.get("VCN_BENCH_GRAPH", "1") == "1":
            self.graphed = V.networks.GraphedSynthesis(self.S, self.n, noise_mode="const")

    def step(self):
        if self.graphed is not None:
            self.graph_launches = getattr(self, "graph_launches", 0) + self.graphed.launches
            return self.graphed(self.ws)
        with torch.no_grad():
            return self.S(self.ws, noise_mode="const")

    def e2e_setup(self, device):
        self.hws = self.ws.cpu().pin_memory()
        self.himg = torch.empty(self.n, 3, self.RES, self.RES).pin_memory()
        self.h2d = self.hws.numel() * 4
        self.d2h = self.himg.numel() * 4

    def e2e_step(self, device):
        with torch.no_grad():
            ws = self.hws.to(device, non_blocking=True)
            img = self.graphed(ws) if self.graphed is not None else self.S(ws, noise_mode="const")
            self.himg.copy_(img, non_blocking=True)

    def roofline(self, device, peaks, iters=20):
        """conv_igemm on the heaviest layer shape b256.conv1 as the pipelines run it: [N,128,256,256] x [128,128,3,3], bf16 NHWC
        in, bf16 NHWC out, fused lrelu epilogue; 2*N*Cout*Cin*9*H*W FLOP."""
        M = self.V.modconv
        n, c, r = self.n, 128, self.RES
        x = torch.randn(n, c, r, r, device=device)
        w = torch.randn(c, c, 3, 3, device=device)
        xh = M.modulate_pack(x, None)
        wt, _ = M.pack_weights(w)
        out = torch.empty(n, c, r, r, device=device, dtype=torch.bfloat16, memory_format=torch.channels_last)
        bias = torch.zeros(c, device=device)
        epi, hold = M.make_epilogue(bias=bias, act="lrelu")
        from viscoin_next_b200 import _lib
        ms = time_kernel_flushed(lambda: M.conv_igemm(xh, wt, _lib.CONV_FPROP, c, 3, 3, out=out, epilogue=epi), device, iters)
        flop = 2.0 * n * c * c * 9 * r * r
        ach = flop / (ms * 1e-3) / 1e12
        return {"kernel": "conv_igemm_kernel FPROP [%d,128,256,256]x[128,128,3,3] bf16 NHWC -> bf16 NHWC" % n, "bound": "tensor",
                "achieved": round(ach, 1), "peak": peaks["bf16"], "unit": "TFLOP/s", "frac": round(ach / peaks["bf16"], 4),
                "peak_source": peaks["src"] + " (burst cuBLAS bf16)", "traffic": None, "alg_flop_per_launch": flop,
                "avg_launch_ms": round(ms, 5)}

    def reference_gpu_setup(self, device, seed):
        """Unmodified reference SynthesisNetwork (oracle/_ref) with its plugins + cuDNN, fp32, eval mode (fused grouped
        convolution branch, networks.py:74-84), same weights as `_build`."""
        from oracle import stage_ref
        import viscoin_next_b200 as V
        ref = stage_ref.load_plugins()
        mine = self._build(torch.device("cpu"), V.networks)
        S = ref.networks.SynthesisNetwork(w_dim=512, img_resolution=self.RES, img_channels=3)
        S.load_state_dict(mine.state_dict(), strict=True)
        self.S = S.to(device).eval().requires_grad_(False)
        g = torch.Generator(device=device).manual_seed(seed)
        self.ws = torch.randn(self.n, self.S.num_ws, 512, device=device, generator=g)
        self.graphed = None
        return "reference SynthesisNetwork (oracle/_ref, unmodified) + reference CUDA plugins (sm_100a JIT) + cuDNN, fp32/TF32, eval mode, eager"

    def cpu_step(self, n_sample, seed=0):
        from oracle import ref_net as RN
        if not hasattr(self, "_cpu_sd"):
            import viscoin_next_b200 as V
            self._cpu_sd = {k: v.detach().cpu() for k, v in self._build(torch.device("cpu"), V.networks).state_dict().items()}
        g = torch.Generator().manual_seed(seed)
        ws = torch.randn(n_sample, 14, 512, generator=g)
        t0 = time.perf_counter()
        with torch.no_grad():
            RN.synthesis_network(ws, self._cpu_sd, self.RES, noise_mode="const")
        return time.perf_counter() - t0


class TrainStep:
    """BASELINE.json configs[3]/[4]: the full VisCoIN ensemble train step (ResNet-50 f, Psi, Theta, generator
    reconstruction and its losses, two Adam updates), batch 8 real + 8 fake images per GPU at 256x256, random-init,
    synthetic CUB-shaped data.  Under torchrun every rank runs its own shard and the 96 M trainable gradients are
    all-reduced (mean) over NCCL while backward is still running (viscoin_next_b200.dp.GradBuckets)."""

    metric = "VisCoIN train steps/s @256^2 b8/GPU"
    unit = "steps/s (x GPUs: one step = 8 real + 8 fake images on each GPU)"
    dtype = "bf16"
    l2_note = "activations per step (GBs) far exceed L2; no flush needed"
    arithmetic_note = ("bf16 operands + bf16 NHWC storage with fp32 accumulation (synthesis tensor-core GEMMs; f / Psi / VGG-16 under "
                       "bf16 autocast), fp32 master weights, losses and Adam; the reference runs fp32 (TF32 convolutions on a GPU)")
    step_flop = None  # set in __init__ (6.3 TFLOP at 8 + 8 images)

    parallelism_note = "dp%d (batch-sharded replicas; one NCCL mean all-reduce of the 96.1 M trainable gradients per step, bucketed and overlapped with backward)"

    def __init__(self, batch):
        self.n = batch
        self.name = ("VisCoIN ensemble train step (ResNet-50 f + Psi + Theta + G reconstruction losses + LPIPS-VGG16, Adam), "
                     "%d real + %d fake images/GPU at 256x256, random-init, i > cd_fid_iteration (all six losses)" % (batch, batch))

        # synthesis 90.24 GF/img x (n fake fwd + 2n fwd + 2n dgrad), ResNet-50 10.68 x (4n fwd + 2n bwd),
        # VGG-16 40.09 x (4n + 2n), Psi 5.49 x 2n x 3  ->  6.3 TFLOP at n = 8
        self.step_flop = (90.24 * 5 * batch + 10.68 * 6 * batch + 40.09 * 6 * batch + 5.49 * 6 * batch) * 1e9

    def units_per_step(self):
        return 1.0

    def setup(self, device, seed):
        import viscoin_next_b200 as V
        self.V = V
        self.tr = V.viscoin_step.VisCoINTrainer(device, seed=0)  # identical replicas (seed 0 + broadcast)
        g = torch.Generator(device=device).manual_seed(seed)     # rank-distinct data
        self.real = torch.randn(self.n, 3, 256, 256, device=device, generator=g)
        self.labels = torch.randint(0, 200, (self.n,), device=device, generator=g)
        self.zgen = torch.Generator().manual_seed(seed)
        self.last = None
        self.graph = os.environ.get("VCN_BENCH_GRAPH", "1") == "1"

    @property
    def graph_launches(self):
        return self.tr.graph_launches

    def _step(self, real, labels):
        if self.graph:  # whole iteration replayed from one CUDA graph (first calls run eagerly and capture)
            return self.tr.step_graphed(real, labels, 1000, z_generator=self.zgen)
        return self.tr.step(real.to(self.tr.device, non_blocking=True), labels.to(self.tr.device, non_blocking=True), 1000,
                            z_generator=self.zgen)

    def step(self):
        self.last = self._step(self.real, self.labels)
        return self.last

    def e2e_setup(self, device):
        self.hreal = self.real.cpu().pin_memory()
        self.hlabels = self.labels.cpu().pin_memory()
        self.h2d = self.hreal.numel() * 4 + self.hlabels.numel() * 8 + self.n * 512 * 4  # images, labels, z of the fake batch
        self.d2h = 4

    def e2e_step(self, device):
        out = self._step(self.hreal, self.hlabels)  # pinned host buffers -> device inside the step
        self.loss_value = float(out["total"])  # device -> host read of the step's result

    def roofline(self, device, peaks, iters=20):
        r = Synthesis.roofline(self, device, peaks, iters)
        r["note"] = "dominant hand-written kernel of the step: the tcgen05 implicit GEMM at the heaviest synthesis layer shape"
        return r

    RES = 256

    def cpu_step(self, n_sample, seed=0):
        """Oracle port of the same step on the host cores: n_sample real + n_sample fake images, forward + backward + BOTH
        Adam updates (viscoin/training/viscoin.py:163-223) - like for like with the GPU step."""
        from oracle import ref_step as RS
        if not hasattr(self, "_cpu_state"):
            import viscoin_next_b200 as V
            VM, N = V.viscoin_models, V.networks
            torch.manual_seed(0)
            sd = lambda m: {k: v.detach().clone() for k, v in m.state_dict().items()}
            gen = N.Generator(z_dim=512, c_dim=0, w_dim=512, img_resolution=256, img_channels=3, mapping_kwargs={"num_layers": 2})
            lp = VM.LPIPS()
            st = {"classifier": sd(VM.Classifier(200)), "concept_extractor": sd(VM.ConceptExtractor()),
                  "explainer": sd(VM.Explainer()), "viscoin_gan": sd(N.GeneratorAdapted.from_gan(gen)),
                  "generator_gan": sd(gen)}
            st["lpips_convs"] = [(l.weight.detach(), l.bias.detach()) for s in lp.slices for l in s if isinstance(l, torch.nn.Conv2d)]
            st["lpips_lins"] = [p.detach() for p in lp.lins]
            st["_f"] = RS.make_classifier(st["classifier"])
            main, gan = [], []
            for grp in ("concept_extractor", "explainer"):
                for v in st[grp].values():
                    if v.is_floating_point():
                        main.append(v.requires_grad_(True))
            for k, v in st["viscoin_gan"].items():
                if k.startswith("mapping."):
                    gan.append(v.requires_grad_(True))
            self._cpu_state = st
            self._cpu_opts = (torch.optim.Adam(main, lr=1e-4), torch.optim.Adam(gan, lr=1e-4))  # viscoin.py:130-134
        st = self._cpu_state
        from oracle import ref_net as RN
        g = torch.Generator().manual_seed(seed)
        real = torch.randn(n_sample, 3, 256, 256, generator=g)
        labels = torch.randint(0, 200, (n_sample,), generator=g)
        t0 = time.perf_counter()
        with torch.no_grad():
            fake = RN.generator(torch.randn(n_sample, 512, generator=g), st["generator_gan"], 256, mapping_layers=2, training=True)
        out = RS.step_losses(st, real, labels, fake, noise_mode="random")
        out.pop("_rebuilt")
        sum(out.values()).backward()
        for opt in self._cpu_opts:
            opt.step()
            opt.zero_grad(set_to_none=True)
        return time.perf_counter() - t0

    # --- the reference's own modules on the GPU (secondary baseline) ------------------------------------
    def reference_gpu_setup(self, device, seed):
        """The UNMODIFIED reference modules (oracle/_ref: Generator / GeneratorAdapted / ConceptExtractor / Explainer with
        their CUDA plugins + cuDNN) in the restated iteration, fp32 parameters, torch's default TF32 convolutions, eager -
        the way `python main.py train viscoin` runs them.  f = HF ResNet-50 and LPIPS-VGG16 (torch formulation) as in the
        reference; none of this repo's kernels is on the path."""
        from oracle import stage_ref
        import viscoin_next_b200 as V
        ref = stage_ref.load_plugins()
        VM = V.viscoin_models
        torch.manual_seed(0)
        cls = VM.Classifier(200)
        psi = ref.concept_extractors.ConceptExtractor()
        theta = ref.explainers.Explainer(n_classes=200)
        G = ref.networks.Generator(z_dim=512, c_dim=0, w_dim=512, img_resolution=256, img_channels=3, mapping_kwargs={"num_layers": 2})
        GA = ref.gan.GeneratorAdapted.from_gan(G)
        lp = VM.LPIPS()
        lp.fused = False
        cls = cls.to(device)
        with torch.no_grad():
            fakes = G.to(device)(torch.randn(4, 512, generator=torch.Generator().manual_seed(0)).to(device), None).float()
        VM.calibrate_batchnorm(cls, torch.cat([torch.randn(4, 3, 256, 256, generator=torch.Generator().manual_seed(1)).to(device), fakes]))
        self.tr = V.viscoin_step.VisCoINTrainer(device, amp=None, channels_last=False, fold_bn=False, fused_adam=False,
                                                models=(cls, psi, theta, G, GA, lp))
        g = torch.Generator(device=device).manual_seed(seed)
        self.real = torch.randn(self.n, 3, 256, 256, device=device, generator=g)
        self.labels = torch.randint(0, 200, (self.n,), device=device, generator=g)
        self.zgen = torch.Generator().manual_seed(seed)
        self.graph = False
        return "reference modules (oracle/_ref, unmodified) + reference CUDA plugins (sm_100a JIT) + cuDNN, fp32 parameters, TF32 convolutions (torch default), eager"


def parity_summary(rep):
    """Worst measured errors of the BENCHMARKED trainer configuration against the CPU oracle (profiles/parity_r2.json,
    written by tests/test_gpu_step.py on a B200)."""
    out = {"source": "profiles/parity_r2.json (tests/test_gpu_step.py on a B200: bf16 autocast + channels_last + folded BN + bf16-resident synthesis vs the fp32 CPU oracle)"}
    for sec in ("bench_l1l2", "bench_all"):
        r = rep.get(sec)
        if not r:
            continue
        groups = {k: v for k, v in r["grads"].items() if "(all" in k}
        out[sec] = {"worst_loss_rel": max(v["rel"] for v in r["losses"].values()),
                    "worst_group_grad_rel_l2": max(v["rel_l2"] for v in groups.values()),
                    "worst_group_grad_cos": min(v["cos"] for v in groups.values()),
                    "worst_adam_update_cos": min((v["update_cos"] for v in r.get("adam", {}).values()), default=None)}
    if "graphed_vs_eager" in rep:
        out["graphed_vs_eager"] = {"worst_loss_rel": rep["graphed_vs_eager"]["worst_loss_rel"],
                                   "params_rel_l2": rep["graphed_vs_eager"]["params"]["rel_l2"],
                                   "bitwise_equal": rep["graphed_vs_eager"]["bitwise_equal"]}
    if "dp" in rep:
        out["dp_2rank_vs_1rank"] = {k: v["rel_l2"] for k, v in rep["dp"]["grads"].items()}
    return out


WORKLOADS = {"opsweep": OpSweep, "synthesis": Synthesis, "trainstep": TrainStep}


def pick_workload(args):
    name = args.workload
    if name == "auto":
        name = DEFAULT_WORKLOAD
    if name not in WORKLOADS:
        raise SystemExit("unknown workload %r (have: %s)" % (name, ", ".join(WORKLOADS)))
    return WORKLOADS[name](args.batch)


DEFAULT_WORKLOAD = "trainstep"


def run_reference(args, world, rank):
    """The reference's own CPU implementation of the path (oracle port of its impl='ref' code) on the host cores: the FULL
    batch of the workload per step (for the train step: forward + backward + both Adam updates), --warmup / --steps honoured
    while the whole run fits ~4 minutes (fewer timed steps otherwise; `steps` reports what ran)."""
    if rank != 0:
        return
    wl = pick_workload(args)
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    t_first = wl.cpu_step(wl.n)
    budget = 240.0
    warm = max(0, min(args.warmup, int(budget * 0.2 / max(t_first, 1e-3))) - 1)
    for _ in range(warm):
        wl.cpu_step(wl.n)
    steps = max(1, min(args.steps, int((budget - (warm + 1) * t_first) / max(t_first, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(steps):
        wl.cpu_step(wl.n)
    sec = (time.perf_counter() - t0) / steps
    value = wl.units_per_step() / sec
    line = {"impl": "reference", "metric": wl.metric, "value": round(value, 4), "unit": wl.unit, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm + 1, "ms_per_step": round(sec * 1e3, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl.name, "batch_per_gpu": wl.n, "device": "host CPU"},
            "cpu_baseline": {"value": round(value, 4), "unit": wl.unit, "cores": cores, "kind": "port", "extrapolated": False,
                             "sample": "the full batch (%d images/GPU) per step%s, oracle/ (torch CPU restatement of the reference impl='ref' path); "
                                       "%d timed steps of %d requested" % (wl.n, ", both Adam updates included" if isinstance(wl, TrainStep) else "", steps, args.steps)},
            "e2e": {"value": round(value, 4), "unit": wl.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_reference_gpu(args, world, rank, local):
    """Secondary baseline: the unmodified reference modules on this B200 - its two CUDA plugins JIT-built for
    sm_100a and cuDNN convolutions - timed like our own arm (CUDA events, W warm-up + K steps).  Rank 0 only."""
    if rank != 0:
        return
    from oracle import stage_ref
    if not stage_ref.available() or not torch.cuda.is_available():
        print(json.dumps({"impl": "reference-gpu", "unavailable": "oracle/_ref not staged (build container: python -m oracle.stage_ref) or no GPU"}), flush=True)
        return
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    wl = pick_workload(args)
    how = wl.reference_gpu_setup(device, seed=1000 * world + rank)
    for _ in range(max(args.warmup, 3)):
        wl.step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        wl.step()
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    value = wl.units_per_step() * args.steps / (ms * 1e-3)
    line = {"impl": "reference-gpu", "metric": wl.metric, "value": round(value, 3), "unit": wl.unit, "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (TF32 convolutions)", "data": "synthetic",
            "config": {"workload": wl.name, "batch_per_gpu": wl.n, "how": how}, "clocks": clocks}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    world, rank, local = dist_setup(args)
    if args.impl == "reference":
        run_reference(args, world, rank)
        return
    if args.impl == "reference-gpu":
        run_reference_gpu(args, world, rank, local)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    import viscoin_next_b200 as V
    lib = V._lib.load()
    V._lib.require_device(local)
    peaks = measured_peaks()
    wl = pick_workload(args)
    wl.setup(device, seed=1000 * world + rank)

    for _ in range(max(args.warmup, 3)):
        wl.step()
    torch.cuda.synchronize()
    barrier(world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = lib.vcn_launch_count() + getattr(wl, "graph_launches", 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        wl.step()
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    launches = lib.vcn_launch_count() + getattr(wl, "graph_launches", 0) - launches0  # eager launches + graph-replayed ones
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(e0.elapsed_time(e1), world, device)
    value = wl.units_per_step() * world * args.steps / (ms * 1e-3)

    # end to end through the public API with host buffers
    wl.e2e_setup(device)
    for _ in range(2):
        wl.e2e_step(device)
    torch.cuda.synchronize()
    barrier(world)
    k_e2e = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    e0.record()
    for _ in range(k_e2e):
        wl.e2e_step(device)
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    ms_e2e = max_over_ranks(ms_e2e, world, device)
    e2e_value = wl.units_per_step() * world * k_e2e / (ms_e2e * 1e-3)

    roof = cpu = others = None
    if rank == 0:
        roof = wl.roofline(device, peaks)
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):  # dram bytes per launch from the committed ncu --set full capture
            try:
                roof["traffic"] = json.load(open(prof)).get(roof["kernel"].split("<")[0].split(" ")[0])
            except Exception:
                pass
        if os.environ.get("VCN_BENCH_KERNELS", "1") == "1":
            # every other hand-written kernel alone at its BASELINE shape (the "upfirdn2d/bias_act HBM GB/s vs peak" half of
            # BASELINE.json's metric) and, for the convolutions, the cuDNN kernel of the same layer on this box
            from tools import kernel_rooflines
            try:
                others = kernel_rooflines.collect(device, peaks, time_kernel_flushed, iters=10, n=wl.n)
            except Exception as e:  # pragma: no cover - a failing side measurement must not lose the headline line
                others = [{"error": "%s: %s" % (type(e).__name__, str(e)[:200])}]
        if world == 1 and not args.no_cpu_baseline:
            cores = len(os.sched_getaffinity(0))
            torch.set_num_threads(cores)
            wl.cpu_step(wl.n)
            sec = wl.cpu_step(wl.n)
            cpu = {"value": round(wl.units_per_step() / sec, 4), "unit": wl.unit, "cores": cores, "kind": "port", "extrapolated": False,
                   "sample": "one full step (%d images/GPU%s) after one warm-up step, oracle/ (torch CPU restatement of the reference impl='ref' path)"
                             % (wl.n, ", forward + backward + both Adam updates" if isinstance(wl, TrainStep) else "")}
    barrier(world)
    if rank == 0:
        line = {"metric": wl.metric, "value": round(value, 3), "unit": wl.unit, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic",
                "config": {"workload": wl.name, "batch_per_gpu": wl.n, "l2": getattr(wl, "l2_note", "working set per step larger than L2 (activations >= 67 MB per layer at 64^2 and above)"),
                           "parallelism": getattr(wl, "parallelism_note", "dp%d (independent shards, no data-path collective)") % world,
                           "arithmetic": getattr(wl, "arithmetic_note", "fp32 storage and accumulation")},
                "clocks": clocks,
                "e2e": {"value": round(e2e_value, 3), "unit": wl.unit, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                        "steps": k_e2e},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu}
        flop = getattr(wl, "step_flop", None)
        if flop:  # whole-step arithmetic rate against the SUSTAINED bf16 peak (the step runs for seconds under the power cap)
            tf = flop / (ms / args.steps * 1e-3) / 1e12
            line["step_tflops"] = round(tf, 1)
            line["step_frac_of_sustained"] = round(tf / peaks["bf16_sus"], 4)
        others = others or getattr(wl, "other_rooflines", None)
        if others:
            line["roofline_others"] = others
        par = os.path.join(ROOT, "profiles", "parity_r2.json")
        if os.path.exists(par):  # measured errors of the benchmarked configuration vs the oracle (tests/test_gpu_step.py)
            try:
                line["parity"] = parity_summary(json.load(open(par)))
            except Exception:
                pass
        print(json.dumps(line), flush=True)
    if world > 1:
        # Orderly teardown: the CUDA graphs that captured NCCL collectives go first (they pin the communicator's streams), then
        # the process group.  A daemon watchdog leaves the process if the communicator still refuses to come down - the line
        # is printed and every rank is synchronised by then, so a stuck teardown must not cost the run.
        sys.stdout.flush()
        sys.stderr.flush()
        torch.cuda.synchronize()
        barrier(world)
        import gc
        import threading
        dog = threading.Timer(20.0, lambda: os._exit(0))
        dog.daemon = True
        dog.start()
        tr = getattr(wl, "tr", None)
        if tr is not None:
            tr._graphs.clear()
        del wl, tr
        gc.collect()
        torch.cuda.synchronize()
        torch.distributed.destroy_process_group()
        dog.cancel()


if __name__ == "__main__":
    main()
