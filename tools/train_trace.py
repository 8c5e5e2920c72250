"""Kernel timeline of the captured training step (CUPTI through torch.profiler, graph replays): per-kernel time per step,
and the main stream's busy time against the step time.  Usage: train_trace.py [N_rays] [K]"""
import os, sys
from types import SimpleNamespace
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
from dm_nerf_b200.train import TrainStep

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
k = int(sys.argv[2]) if len(sys.argv) > 2 else 32
near, far = (4.0, 15.0) if k == 32 else (0.0, 6.5)
ops.set_mode("bf16")
H, W = 480, 640
dev = torch.device("cuda")
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
Kmat, c2w = synth.dmsr_camera(H, W, 0)
pose = c2w[:3, :4].to(dev)
img = torch.rand(H, W, 3, device=dev)
lab = synth.block_labels(H, W, k, device=dev)
targs = SimpleNamespace(N_train=n, near=near, far=far, N_samples=64, N_importance=128, perturb=1.0, ins_num=k,
                        tolerance=0.05, deta_w=0.05, lrate=5e-4, lrate_decay=500)
mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
mc.load_state_dict(synth.init_state_dict(k, 1)); mf.load_state_dict(synth.init_state_dict(k, 2))
ts = TrainStep(pe, ve, mc.to(dev), mf.to(dev), Kmat, targs, penalize=True, graph=True)
for _ in range(10):
    ts.step(img, pose, lab)
torch.cuda.synchronize()
steps = 10
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(steps):
    ts.step(img, pose, lab)
b.record(); torch.cuda.synchronize()
print(f"N={n} K={k}: {a.elapsed_time(b) / steps:.3f} ms/step (events)")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(steps):
        ts.step(img, pose, lab)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
by = {}
streams = {}
for e in evs:
    nm = e.name[:70]
    t = by.setdefault(nm, [0, 0.0]); t[0] += 1; t[1] += e.device_time
tot = sum(v[1] for v in by.values())
print(f"sum of kernel time over all streams: {tot / steps / 1000:.3f} ms/step, {sum(v[0] for v in by.values()) / steps:.0f} kernels/step")
for nm, (c, t) in sorted(by.items(), key=lambda x: -x[1][1])[:28]:
    print(f"  {t / steps:9.1f} us  x{c / steps:5.1f}  {nm}")

# ---- per-stream timeline of the LAST profiled step: gaps on the busiest (main) stream ----
kin = [e for e in prof.profiler.kineto_results.events() if e.device_type() == torch.autograd.DeviceType.CUDA and e.duration_ns() > 0]
kin.sort(key=lambda e: e.start_ns())
by_stream = {}
for e in kin:
    by_stream.setdefault(e.device_resource_id(), []).append(e)
main_id = max(by_stream, key=lambda s: sum(e.duration_ns() for e in by_stream[s]))
main = by_stream[main_id]
t_end = main[-1].start_ns() + main[-1].duration_ns()
span = a.elapsed_time(b) / steps * 1e6        # ns per step
last = [e for e in main if e.start_ns() >= t_end - span]
print(f"main stream {main_id}: last step has {len(last)} kernels, busy {sum(e.duration_ns() for e in last) / 1e3:.0f} us of {span / 1e3:.0f} us")
prev_end = None
for e in last:
    if prev_end is not None and e.start_ns() - prev_end > 15000:
        print(f"  gap {(e.start_ns() - prev_end) / 1e3:7.1f} us before {e.name()[:60]}  (at +{(e.start_ns() - last[0].start_ns()) / 1e3:.0f} us)")
    prev_end = e.start_ns() + e.duration_ns()
for sid, evs in by_stream.items():
    if sid == main_id: continue
    ev = [e for e in evs if e.start_ns() >= t_end - span]
    if ev:
        print(f"  stream {sid}: {len(ev)} kernels, busy {sum(e.duration_ns() for e in ev) / 1e3:.0f} us, from +{(ev[0].start_ns() - last[0].start_ns()) / 1e3:.0f} to +{(ev[-1].start_ns() + ev[-1].duration_ns() - last[0].start_ns()) / 1e3:.0f} us")

# ---- what surrounds the largest main-stream gap ----
gaps = []
for i in range(1, len(last)):
    pe_ = last[i - 1].start_ns() + last[i - 1].duration_ns()
    gaps.append((last[i].start_ns() - pe_, i))
g, gi = max(gaps)
t0 = last[0].start_ns()
print(f"largest gap {g / 1e3:.0f} us; main stream around it:")
for e in last[max(0, gi - 5):gi + 4]:
    print(f"   +{(e.start_ns() - t0) / 1e3:7.0f} us {e.duration_ns() / 1e3:7.1f} us  {e.name()[:80]}")
lo, hi = last[gi - 1].start_ns() + last[gi - 1].duration_ns(), last[gi].start_ns()
for sid, evs in by_stream.items():
    if sid == main_id: continue
    ev = [e for e in evs if e.start_ns() + e.duration_ns() > lo - 50000 and e.start_ns() < hi + 50000]
    print(f"  stream {sid} during the gap ({len(ev)} kernels):")
    for e in ev[:6] + ev[-10:]:
        print(f"   +{(e.start_ns() - t0) / 1e3:7.0f} us {e.duration_ns() / 1e3:7.1f} us  {e.name()[:80]}")

# ---- the heavy kernels of the last step in time order, with the idle time (no kernel on ANY stream) in front of each ----
allk = [e for e in kin if e.start_ns() >= t_end - span]
allk.sort(key=lambda e: e.start_ns())
cover_end = allk[0].start_ns()
idle_total = 0
marks = []
for e in allk:
    idle = max(0, e.start_ns() - cover_end)
    idle_total += idle
    if e.duration_ns() > 7000 or idle > 5000:
        marks.append((e.start_ns(), e.duration_ns(), idle, e.name()[:64], e.device_resource_id()))
    cover_end = max(cover_end, e.start_ns() + e.duration_ns())
print(f"all streams: {len(allk)} kernels, idle (no kernel anywhere) {idle_total / 1e3:.0f} us of {span / 1e3:.0f} us")
for st, du, idle, nm, sid in marks:
    print(f"   +{(st - allk[0].start_ns()) / 1e3:7.0f} us {du / 1e3:7.1f} us  idle-before {idle / 1e3:6.1f}  s{sid:<4d} {nm}")
