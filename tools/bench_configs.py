"""The BASELINE.json configs that are not bench.py's headline line, timed on one GPU:
  C4  ScanNet-shaped 1296x968 full-frame instance render, K=16, near 0 / far 9.5 (one rank's view of the frame;
      with G ranks each renders 1/G of the rows, frame.render_pixels(pix0, count))
  C5  manipulation demo view: 480x640, K=32, one rigidly moved object (T=1), random uniforms as the reference draws them
Prints one JSON line per config (device time, CUDA events, 3 warm-ups)."""
import json, math, os, sys
from types import SimpleNamespace
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import ops
from dm_nerf_b200.frame import render_pixels, manipulate_pixels
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder

ops.set_mode("bf16")
dev = torch.device("cuda")
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3


def models(k):
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
    return mc.to(dev), mf.to(dev)


def timed(fn, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


# ---- C4 ----
H, W, k = 968, 1296, 16
mc, mf = models(k)
Kmat = torch.tensor([[1170.0, 0, 648.0], [0, 1170.0, 484.0], [0, 0, 1]])
pose = orc.pose_spherical(37.0, -65.0, 4.0)
out = None
def c4():
    global out
    out = render_pixels(pe, ve, mc, mf, (H, W, Kmat), pose, 0.0, 9.5, 64, 128, chunk=32768, out=out)
ms = timed(c4)
flop = 2.0 * (691840 + 128 * k) * 256 * H * W
print(json.dumps({"config": "C4 ScanNet-shaped 1296x968 render, K=16, 64+128 samples, 1 GPU", "ms_per_frame": ms,
                  "rays_per_s": H * W / ms * 1e3, "algorithmic_tflops": flop / ms / 1e9}))

# ---- C5 ----
H, W, k = 480, 640, 32
mc, mf = models(k)
focal = 0.5 * W / math.tan(0.5 * 0.6911)
Kmat = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]])
pose = orc.pose_spherical(37.0, -65.0, 7.0)
th = math.radians(30.0)
c = torch.tensor([-0.52, -2.17, 1.07])
R = torch.tensor([[math.cos(th), -math.sin(th), 0], [math.sin(th), math.cos(th), 0], [0, 0, 1.0]])
trans = torch.eye(4); trans[:3, :3] = R; trans[:3, 3] = c - R @ c            # rotation about the object centre
args = SimpleNamespace(N_samples=64, N_importance=128, near=4.0, far=15.0)
out5 = None
def c5():
    global out5
    out5 = manipulate_pixels(pe, ve, mc, mf, (H, W, Kmat), pose, [{"tar_id": 8, "transformation": trans}], args,
                             chunk=16384, out=out5)
ms = timed(c5, warm=1)
pts = 64 + 64 + 192 + 192 + 320 + 320                                        # MLP points per ray at T=1
flop = 2.0 * (691840 + 128 * k) * pts * H * W
print(json.dumps({"config": "C5 manipulation view 480x640, K=32, T=1 (1152 MLP points/ray)", "ms_per_view": ms,
                  "rays_per_s": H * W / ms * 1e3, "algorithmic_tflops": flop / ms / 1e9,
                  "sweep_360_views_s": ms * 360 / 1e3}))
