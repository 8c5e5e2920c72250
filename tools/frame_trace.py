"""Kernel timeline of one rendered frame / one manipulated view (CUPTI through torch.profiler): GPU idle time between
kernels and the non-MLP kernels by total time.  Usage: frame_trace.py [render|manip] [mode]"""
import os, sys, math
from types import SimpleNamespace
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.frame import render_pixels, manipulate_pixels
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
what = sys.argv[1] if len(sys.argv) > 1 else "render"
mode = sys.argv[2] if len(sys.argv) > 2 else "fp16"
H, W, K = 480, 640, 32
dev = "cuda"
mc, mf = DM_NeRF(8, 256, 63, 27, [4], K), DM_NeRF(8, 256, 63, 27, [4], K)
mc.load_state_dict(synth.init_state_dict(K, 1)); mf.load_state_dict(synth.init_state_dict(K, 2))
mc, mf = mc.to(dev), mf.to(dev)
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
Kmat, c2w = synth.dmsr_camera(H, W, 0)
ops.set_mode(mode)
ang = math.radians(30.0)
rot = torch.tensor([[math.cos(ang), -math.sin(ang), 0, 0], [math.sin(ang), math.cos(ang), 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
margs = SimpleNamespace(N_samples=64, N_importance=128, near=4.0, far=15.0)
def run():
    if what == "render":
        render_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, 4.0, 15.0)
    else:
        manipulate_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, [{"tar_id": 8, "transformation": rot}], margs)
for _ in range(2): run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); run(); b.record(); torch.cuda.synchronize()
print(f"{what} {mode}: {a.elapsed_time(b):.2f} ms")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    run(); torch.cuda.synchronize()
kin = [e for e in prof.profiler.kineto_results.events() if e.device_type() == torch.autograd.DeviceType.CUDA and e.duration_ns() > 0]
kin.sort(key=lambda e: e.start_ns())
end = kin[0].start_ns(); idle = 0; gaps = []
by = {}
for e in kin:
    g = max(0, e.start_ns() - end)
    idle += g
    if g > 20000: gaps.append((g, e.name()[:60]))
    end = max(end, e.start_ns() + e.duration_ns())
    t = by.setdefault(e.name()[:70], [0, 0]); t[0] += 1; t[1] += e.duration_ns()
span = end - kin[0].start_ns()
print(f"{len(kin)} kernels, span {span / 1e6:.2f} ms, idle {idle / 1e3:.0f} us ({100 * idle / span:.2f} %), {len(gaps)} gaps > 20 us")
for g, nm in sorted(gaps, reverse=True)[:8]:
    print(f"   gap {g / 1e3:.0f} us before {nm}")
for nm, (c, t) in sorted(by.items(), key=lambda x: -x[1][1])[:16]:
    print(f"  {t / 1e3:9.1f} us  x{c:4d}  {nm}")
