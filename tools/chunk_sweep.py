"""Rays per pass of the 640x480 frame (C2): ms per frame through dmnerf_render_full for several chunk sizes.
usage: python tools/chunk_sweep.py [mode] [frames] [chunk ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.frame import render_pixels
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder

mode = sys.argv[1] if len(sys.argv) > 1 else "fp16"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 10
chunks = [int(c) for c in sys.argv[3:]] or [16384, 32768, 65536, 102400, 153600, 307200]
H, W, k = 480, 640, 32
ops.set_mode(mode)
mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
mc.load_state_dict(synth.init_state_dict(k, 1)); mf.load_state_dict(synth.init_state_dict(k, 2))
mc, mf = mc.cuda(), mf.cuda()
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
Kmat, c2w = synth.dmsr_camera(H, W, 0)
out = None
for rep in range(2):
    for chunk in chunks:
        for _ in range(2):
            out = render_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, 4.0, 15.0, chunk=chunk, out=out)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(frames):
            out = render_pixels(pe, ve, mc, mf, (H, W, Kmat), c2w, 4.0, 15.0, chunk=chunk, out=out)
        b.record(); torch.cuda.synchronize()
        print(f"chunk {chunk:7d}: {a.elapsed_time(b) / frames:.3f} ms/frame", flush=True)
