"""One 16,384-ray chunk of the manipulated view (C5: mani_demo, T=1) bracketed by cudaProfilerStart/Stop, for
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:exchange -o gpurun_out/r02_exchange python tools/profile_manip.py"""
import os, sys, math
from types import SimpleNamespace
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.frame import manipulate_pixels
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder

mode = sys.argv[1] if len(sys.argv) > 1 else "fp16"
H, W, K = 480, 640, 32
dev = "cuda"
mc, mf = DM_NeRF(8, 256, 63, 27, [4], K), DM_NeRF(8, 256, 63, 27, [4], K)
mc.load_state_dict(synth.init_state_dict(K, 1)); mf.load_state_dict(synth.init_state_dict(K, 2))
mc, mf = mc.to(dev), mf.to(dev)
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
Kmat, c2w = synth.dmsr_camera(H, W, 0)
ang = math.radians(30.0)
rot = torch.tensor([[math.cos(ang), -math.sin(ang), 0, 0], [math.sin(ang), math.cos(ang), 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
targets = [{"tar_id": 8, "transformation": rot}]
args = SimpleNamespace(N_samples=64, N_importance=128, near=4.0, far=15.0)
ops.set_mode(mode)
torch.manual_seed(0)
for _ in range(2):
    manipulate_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, targets, args, pix0=0, count=16384)
torch.cuda.synchronize()
torch.cuda.profiler.start()
manipulate_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, targets, args, pix0=0, count=16384)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
