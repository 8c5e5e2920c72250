"""Ray / sampling helpers with the reference's names (networks/helpers.py)."""
from __future__ import annotations

import numpy as np
import torch

from .. import ops


def get_rays_k(H, W, K, c2w):
    """helpers.py:50-61: rays_o, rays_d [H,W,3] from intrinsics K and camera-to-world c2w."""
    dev = c2w.device if isinstance(c2w, torch.Tensor) and c2w.is_cuda else torch.device("cuda", torch.cuda.current_device())
    ro, rd = ops.rays_k(int(H), int(W), K, c2w, 0, int(H) * int(W), dev)
    return ro.reshape(int(H), int(W), 3), rd.reshape(int(H), int(W), 3)


def get_rays(H, W, focal, c2w):
    """helpers.py:24-35 expressed through the K form: dirs = [(i-(W-1)/2)/f, (j-(H-1)/2)/f, 1]."""
    K = torch.tensor([[focal, 0, (W - 1) * .5], [0, focal, (H - 1) * .5], [0, 0, 1]], dtype=torch.float32)
    return get_rays_k(H, W, K, c2w)


def get_select_full(rgb, pose, K, ins_target, N_train, numpy_rng=False):
    """helpers.py:99-111: N_train distinct random pixels, their rays (generated for those pixels only, not the
    whole frame), colours and instance labels.  The pixel draw runs on the device (``torch.randperm``); the
    reference's ``np.random.choice(H*W, N, replace=False)`` permutes all H*W indices on the host every step
    (3.7 ms at 640x480) -- pass ``numpy_rng=True`` to reproduce its RNG stream."""
    H, W, _ = rgb.shape
    if numpy_rng:
        idx = torch.from_numpy(np.random.choice(H * W, size=[N_train], replace=False)).to(rgb.device)
    else:
        idx = torch.randperm(H * W, device=rgb.device)[:N_train]
    if torch.is_tensor(K) and K.is_cuda and torch.is_tensor(pose) and pose.is_cuda:
        ro, rd = ops.rays_k_dev(int(H), int(W), K, pose, int(N_train), idx)          # camera stays on the device
    else:
        ro, rd = ops.rays_k(int(H), int(W), K, pose, 0, int(N_train), rgb.device, pix_idx=idx)
    batch_rays = torch.stack([ro, rd], 0)
    return rgb.reshape(-1, rgb.shape[-1])[idx], ins_target.reshape(-1)[idx], batch_rays


def z_val_sample(N_rays, near, far, N_samples):
    """helpers.py:114-119: near + linspace(0,1,S)*(far-near), one row broadcast to every ray."""
    dev = torch.device("cuda", torch.cuda.current_device())
    lo = near * torch.ones(size=(N_rays, 1), device=dev)
    hi = far * torch.ones(size=(N_rays, 1), device=dev)
    t = torch.linspace(0., 1., steps=N_samples, device=dev)
    return (lo + t * (hi - lo)).expand([N_rays, N_samples])


def sample_pdf(bins, weights, N_samples, det=False, u=None, cdf=None):
    """helpers.py:123-155.  Uniforms are drawn here with torch.rand in the reference's order
    (or pass ``u``); ``cdf`` switches to the strict mode that takes a torch-computed CDF."""
    lead = bins.shape[:-1]
    bins2 = bins.reshape(-1, bins.shape[-1])
    w2 = weights.reshape(-1, weights.shape[-1])
    if det:
        u = torch.linspace(0., 1., steps=N_samples, device=bins.device)
    elif u is None:
        u = torch.rand(list(lead) + [N_samples], device=bins.device)
    if u.dim() > 1:
        u = u.reshape(-1, N_samples)
    out = ops.sample_pdf(bins2, w2, int(N_samples), u, cdf=None if cdf is None else cdf.reshape(-1, cdf.shape[-1]))
    return out.reshape(*lead, N_samples)
