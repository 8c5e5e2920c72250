"""One render pass of the benchmarked frame (30,720 rays = a tenth of the 640x480 DM-SR frame: coarse MLP, compositing,
fine_z, fine MLP, compositing) bracketed by cudaProfilerStart/Stop, for
  ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/r02_chunk python tools/profile_chunk.py [mode]
mode: fp16 (default) | bf16 | tc32 | fp32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.frame import render_pixels
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder

mode = sys.argv[1] if len(sys.argv) > 1 else "fp16"
count = int(sys.argv[2]) if len(sys.argv) > 2 else 30720
H, W, K = 480, 640, 32
dev = "cuda"
mc, mf = DM_NeRF(8, 256, 63, 27, [4], K), DM_NeRF(8, 256, 63, 27, [4], K)
mc.load_state_dict(synth.init_state_dict(K, 1)); mf.load_state_dict(synth.init_state_dict(K, 2))
mc, mf = mc.to(dev), mf.to(dev)
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
Kmat, c2w = synth.dmsr_camera(H, W, 0)
ops.set_mode(mode)
for _ in range(2):
    render_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, 4.0, 15.0, pix0=0, count=count)
torch.cuda.synchronize()
torch.cuda.profiler.start()
render_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, 4.0, 15.0, pix0=0, count=count)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", mode, count)
