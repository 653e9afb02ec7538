"""Run a few VisCoIN train steps on cuda:0 and print the time per step and a per-phase breakdown (CUDA events).

    python tools/run_step.py [steps] [amp: bf16|fp32]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import viscoin_next_b200 as V  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
amp = torch.bfloat16 if (len(sys.argv) < 3 or sys.argv[2] == "bf16") else None
dev = torch.device("cuda:0")
t0 = time.time()
tr = V.viscoin_step.VisCoINTrainer(dev, amp=amp)
print("built in %.1fs, trainable %.1f M params, %d buckets" % (time.time() - t0, tr.trainable_numel() / 1e6, len(tr.buckets.buckets)), flush=True)
g = torch.Generator(device=dev).manual_seed(0)
real = torch.randn(8, 3, 256, 256, device=dev, generator=g)
labels = torch.randint(0, 200, (8,), device=dev, generator=g)


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


for it in range(steps):
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    e = [ev()]
    losses = tr.forward_losses(real, labels, 1000)
    e.append(ev())
    total = sum(losses.values())
    total.backward()
    e.append(ev())
    tr.buckets.finish()
    tr.optimizer.step()
    tr.gan_optimizer.step()
    tr.buckets.reset()
    e.append(ev())
    torch.cuda.synchronize()
    w1 = time.perf_counter()
    print("step %d: total loss %.4f | fwd %.2f ms  bwd %.2f ms  opt %.2f ms | gpu %.2f ms  wall %.2f ms" % (
        it, float(total), e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[2].elapsed_time(e[3]), e[0].elapsed_time(e[3]),
        (w1 - w0) * 1e3), flush=True)
print({k: round(float(v), 5) for k, v in losses.items()})
if os.environ.get("VCN_STEP_GRAPH", "1") == "1":
    del tr
    torch.cuda.empty_cache()
    tr = V.viscoin_step.VisCoINTrainer(dev, amp=amp)  # fresh: no gradient accumulator may be bound to the default stream
    zg = torch.Generator().manual_seed(1)
    for it in range(8):
        torch.cuda.synchronize()
        w0 = time.perf_counter()
        e0 = ev()
        out = tr.step_graphed(real, labels, 1000, z_generator=zg)
        e1 = ev()
        torch.cuda.synchronize()
        print("graphed step %d: total %.4f gpu %.2f ms wall %.2f ms" % (it, float(out["total"]), e0.elapsed_time(e1), (time.perf_counter() - w0) * 1e3), flush=True)
print("max mem %.2f GB" % (torch.cuda.max_memory_allocated() / 2 ** 30))

if os.environ.get("VCN_STEP_PROFILE", "0") == "1":
    tr = V.viscoin_step.VisCoINTrainer(dev, amp=amp)
    tr.step(real, labels)
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        tr.step(real, labels)
        torch.cuda.synchronize()
    from torch.autograd import DeviceType
    cats, tot = {}, 0.0
    for e in prof.key_averages():
        if e.device_type != DeviceType.CUDA:
            continue
        name, t = e.key, e.self_device_time_total
        if "vcn::" in name:
            c = "ours  " + name.split("::")[-1].split("(")[0].split("<")[0]
        elif "elementwise" in name or "reduce_kernel" in name or "vectorized" in name or "Pool" in name or "pool" in name:
            c = "torch elementwise/reduce/pool"
        elif "batch_norm" in name or "bn_" in name:
            c = "torch batch_norm"
        elif "Memcpy" in name or "Memset" in name:
            c = "memcpy/memset"
        elif "adam" in name.lower():
            c = "fused adam"
        else:
            c = "library conv/gemm + other"
            cats.setdefault("  lib: " + name[:60], [0.0, 0])
            cats["  lib: " + name[:60]][0] += t
            cats["  lib: " + name[:60]][1] += e.count
        cats.setdefault(c, [0.0, 0])
        cats[c][0] += t
        cats[c][1] += e.count
        tot += t
    top = sorted(cats.items(), key=lambda kv: -kv[1][0])
    for k, (t, n) in [kv for kv in top if not kv[0].startswith("  lib")] + [kv for kv in top if kv[0].startswith("  lib")][:12]:
        print("%9.3f ms %5d x  %s" % (t / 1e3, n, k))
    print("device total %.3f ms" % (tot / 1e3))
