"""Host-side breakdown of a training step (wall time per section, with a device sync after each)."""
import os, sys, time
from types import SimpleNamespace
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
from dm_nerf_b200.networks.render import dm_nerf
from dm_nerf_b200.networks.helpers import z_val_sample, get_select_full
from dm_nerf_b200.networks.penalizer import ins_penalizer
from dm_nerf_b200.networks.evaluator import img2mse, ins_criterion

n, k = int(sys.argv[1]), int(sys.argv[2])
ops.set_mode("bf16")
H, W = 480, 640
dev = torch.device("cuda")
mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
mc, mf = mc.to(dev), mf.to(dev)
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=5e-4)
focal = 0.5 * W / np.tan(0.5 * 0.6911)
Kmat = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
pose = torch.eye(4)[:3, :4].clone(); pose[2, 3] = 7.0; pose = pose.to(dev)
img = torch.rand(H, W, 3, device=dev)
lab = ((torch.arange(H, device=dev)[:, None] // 60) * 8 + (torch.arange(W, device=dev)[None] // 80)) % k
args = SimpleNamespace(perturb=1.0, N_importance=128, is_train=True, N_ins=None, tolerance=0.05, deta_w=0.05)
zc = z_val_sample(n, 4.0, 15.0, 64)
T = {}
def tick(name, t0):
    torch.cuda.synchronize(); T[name] = T.get(name, 0.0) + time.perf_counter() - t0; return time.perf_counter()
for it in range(8):
    if it == 3: T.clear()
    t = time.perf_counter()
    target_c, target_i, rays = get_select_full(img, pose, Kmat, lab, n); t = tick("select", t)
    info = dm_nerf(rays, pe, ve, mc, mf, zc, args); t = tick("dm_nerf fwd", t)
    loss = img2mse(info["rgb_coarse"], target_c) + img2mse(info["rgb_fine"], target_c); t = tick("mse", t)
    loss = loss + ins_criterion(info["ins_coarse"], target_i, k)[0] + ins_criterion(info["ins_fine"], target_i, k)[0]; t = tick("ins_criterion x2", t)
    loss = loss + ins_penalizer(info["raw_coarse"], info["z_vals_coarse"], info["depth_coarse"], rays[1], args) \
        + ins_penalizer(info["raw_fine"], info["z_vals_fine"], info["depth_fine"], rays[1], args); t = tick("penalizer x2", t)
    opt.zero_grad(); t = tick("zero_grad", t)
    loss.backward(); t = tick("backward", t)
    opt.step(); t = tick("adam", t)
for kname, v in T.items():
    print(f"{kname:20s} {v / 5 * 1e3:7.3f} ms")
print("total", sum(T.values()) / 5 * 1e3)
