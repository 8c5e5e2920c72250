"""Times one DM-NeRF training step (train_dmsr.py:24-72) on the CUDA path: ray selection, dm_nerf (coarse+fine),
MSE + Hungarian instance loss + emptiness penalizer, backward, Adam.  Usage: train_step.py [N_rays] [K] [steps]"""
import os
import sys
import time
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc  # noqa: E402
from dm_nerf_b200 import ops  # noqa: E402
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder  # noqa: E402
from dm_nerf_b200.networks.render import dm_nerf  # noqa: E402
from dm_nerf_b200.networks.helpers import z_val_sample, get_select_full  # noqa: E402
from dm_nerf_b200.networks.penalizer import ins_penalizer  # noqa: E402
from dm_nerf_b200.networks.evaluator import img2mse, ins_criterion  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    ops.set_mode(sys.argv[4] if len(sys.argv) > 4 else "fp32")
    H, W = 480, 640
    dev = torch.device("cuda")
    np.random.seed(0); torch.manual_seed(0)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
    mc, mf = mc.to(dev), mf.to(dev)
    pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
    opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=5e-4, betas=(0.9, 0.999))
    focal = 0.5 * W / np.tan(0.5 * 0.6911)
    Kmat = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
    pose = torch.eye(4)[:3, :4].clone(); pose[2, 3] = 7.0
    img = torch.rand(H, W, 3, device=dev)
    lab = ((torch.arange(H, device=dev)[:, None] // 60) * 8 + (torch.arange(W, device=dev)[None] // 80)) % k
    args = SimpleNamespace(perturb=1.0, N_importance=128, is_train=True, N_ins=None, tolerance=0.05, deta_w=0.05)
    zc = z_val_sample(n, 4.0, 15.0, 64)

    def step():
        target_c, target_i, rays = get_select_full(img, pose.to(dev), Kmat, lab, n)
        info = dm_nerf(rays, pe, ve, mc, mf, zc, args)
        loss = img2mse(info["rgb_coarse"], target_c) + img2mse(info["rgb_fine"], target_c)
        loss = loss + ins_criterion(info["ins_coarse"], target_i, k)[0] + ins_criterion(info["ins_fine"], target_i, k)[0]
        loss = loss + ins_penalizer(info["raw_coarse"], info["z_vals_coarse"], info["depth_coarse"], rays[1], args) \
            + ins_penalizer(info["raw_fine"], info["z_vals_fine"], info["depth_fine"], rays[1], args)
        opt.zero_grad()
        loss.backward()
        opt.step()
        return loss

    for _ in range(2):
        l0 = step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        l = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    print(f"train step N={n} K={k} mode={ops.get_mode()}: {dt * 1e3:.2f} ms/step -> {n / dt:.0f} rays/s; loss {float(l0.detach()):.4f} -> {float(l.detach()):.4f}")
    if os.environ.get("DMNERF_NO_TORCH_PROFILER"):
        return
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step(); torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))


if __name__ == "__main__":
    main()
