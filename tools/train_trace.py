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
