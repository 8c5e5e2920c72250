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


def get_select_full(rgb, pose, K, ins_target, N_train, numpy_rng=False, idx=None):
    """helpers.py:99-111: N_train distinct random pixels, their rays (generated for those pixels only, not the
    whole frame), colours and instance labels.  The pixel draw runs on the device (``torch.randperm``); the
    reference's ``np.random.choice(H*W, N, replace=False)`` permutes all H*W indices on the host every step
    (3.7 ms at 640x480) -- pass ``numpy_rng=True`` to reproduce its RNG stream."""
    H, W, _ = rgb.shape
    if idx is not None:
        pass                                            # drawn by the caller (ops.draw_pixels in the captured training step)
    elif numpy_rng:
        idx = torch.from_numpy(np.random.choice(H * W, size=[N_train], replace=False)).to(rgb.device)
    else:
        idx = torch.randperm(H * W, device=rgb.device)[:N_train]
    if torch.is_tensor(K) and K.is_cuda and torch.is_tensor(pose) and pose.is_cuda:
        ro, rd = ops.rays_k_dev(int(H), int(W), K, pose, int(N_train), idx)          # camera stays on the device
    else:
        ro, rd = ops.rays_k(int(H), int(W), K, pose, 0, int(N_train), rgb.device, pix_idx=idx)
    batch_rays = torch.stack([ro, rd], 0)
    return rgb.reshape(-1, rgb.shape[-1])[idx], ins_target.reshape(-1)[idx], batch_rays


def get_select_crop(rgb, pose, K, ins_target, ins_index, crop_mask, N_train):
    """helpers.py:64-96 (ScanNet sampling, train_scannet.py): 30 % of the batch from the labelled pixels ``ins_index``,
    the rest from the crop mask; labelled rays come LAST (``dm_nerf`` slices ``[-N_ins:]``, render.py:88-90).
    Host glue with the reference's numpy draws -- including its indexing of ``crop_indices`` (not of the set difference)
    by the unlabelled draw, helpers.py:84-85 -- but rays are generated for the selected pixels only.
    Returns (target_c [N_train,3], target_i [N_ins], batch_rays [2,N_train,3], N_ins)."""
    N_ins = min(int(N_train * 0.3), len(ins_index))
    N_rgb = N_train - N_ins
    H, W, _ = rgb.shape
    crop_indices = np.where(np.asarray(crop_mask.cpu() if torch.is_tensor(crop_mask) else crop_mask).reshape(-1) == 1)[0]
    ins_index_np = np.asarray(ins_index.cpu() if torch.is_tensor(ins_index) else ins_index).reshape(-1)
    labeled_idx = ins_index_np[np.random.choice(ins_index_np.shape[0], size=[N_ins], replace=False)]
    n_unlabeled = len(set(crop_indices.tolist()) - set(labeled_idx.tolist()))
    unlabeled_idx = crop_indices[np.random.choice(n_unlabeled, size=[N_rgb], replace=False)]
    idx = torch.from_numpy(np.concatenate([unlabeled_idx, labeled_idx]).astype(np.int64)).to(rgb.device)
    ro, rd = ops.rays_k(int(H), int(W), K, pose, 0, int(N_train), rgb.device, pix_idx=idx)
    target_c = rgb.reshape(-1, rgb.shape[-1])[idx]
    target_i = ins_target.reshape(-1)[idx[N_rgb:]]
    return target_c, target_i, torch.stack([ro, rd], 0), N_ins


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
