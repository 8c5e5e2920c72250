"""Manipulator core with the reference's surface (networks/manipulator.py:18-205)."""
from __future__ import annotations

import torch

from .. import ops
from .render import _query


def exchanger(ori_raw, tar_raws, ori_raw_pred, tar_raw_preds, move_labels, want_labels: bool = True):
    """manipulator.py:18-83.  Edits ``ori_raw`` in place, one kernel launch per move label (the
    reference's label-order dependence is carried through the running per-sample label tensor).
    ``want_labels=False`` (``manipulator``, which drops them: :177, :201): the int64 label maps are only kept where a
    later move label needs them, the returned ones are None."""
    if isinstance(move_labels, int):
        move_labels = [move_labels]
    ori_label, tar_label = None, None
    last = len(move_labels) - 1
    for idx, m in enumerate(move_labels):
        ori_label, tar_label = ops.exchange(ori_raw, tar_raws[idx], ori_raw_pred, tar_raw_preds[idx], int(m), ori_label,
                                            want_ori_label=want_labels or idx < last, want_tar_label=want_labels)
    return ori_raw, tar_raws, ori_label, tar_label


def manipulator_render(raw, z_vals, rays_d):
    """manipulator.py:86-105 -> (rgb, weights, depth, ins [N,K+1]): keeps the 'empty' channel."""
    return ops.composite(raw, z_vals, rays_d, 0, False)


def _lerp_z(n_rays, near, far, n_samples, device):
    """manipulator.py:116-120: near*(1-t) + far*t (differs from z_val_sample in the last ulp)."""
    lo = near * torch.ones(size=(n_rays, 1), device=device)
    hi = far * torch.ones(size=(n_rays, 1), device=device)
    t = torch.linspace(0., 1., steps=n_samples, device=device)
    return (lo * (1. - t) + hi * t).expand([n_rays, n_samples]).contiguous()


def manipulator_nerf(rays, position_embedder, view_embedder, model, N_samples=None, near=None, far=None, z_vals=None):
    """manipulator.py:108-134 -> (raw [N,S,C], z_vals [N,S])."""
    rays_o, rays_d = rays
    rays_o, rays_d = rays_o.float().contiguous(), rays_d.float().contiguous()
    if z_vals is None:
        z_vals = _lerp_z(rays_d.shape[0], near, far, N_samples, rays_d.device)
    raw = _query(model, rays_o, rays_d, z_vals.contiguous(), position_embedder, view_embedder)
    return raw, z_vals


def manipulator(position_embedder, view_embedder, model_coarse, model_fine, ori_rays, f_tar_rays, args, us=None):
    """manipulator.py:137-205.  ``us``: optional list of uniforms [N,N_importance] in the reference's
    draw order (ori stage 1, one per target, ori stage 2); drawn with torch.rand when absent.
    Accepts a single target as [2,N,3] and an int target_label (the stale manipulator_eval caller)."""
    N_samples, N_importance, near, far = args.N_samples, args.N_importance, args.near, args.far
    if isinstance(f_tar_rays, torch.Tensor) and f_tar_rays.dim() == 3:
        f_tar_rays = f_tar_rays[None]
    labels = args.target_label if isinstance(args.target_label, (list, tuple)) else [args.target_label]
    n = ori_rays[1].shape[0]
    dev = ori_rays[1].device
    us = list(us) if us is not None else None
    draw = (lambda: us.pop(0)) if us is not None else (lambda: torch.rand([n, N_importance], device=dev))
    pe, ve = position_embedder, view_embedder

    def stage1(rays):
        raw, z = manipulator_nerf(rays, pe, ve, model_coarse, N_samples, near, far)
        rgb, w, _, _ = manipulator_render(raw, z, rays[1])
        z_full, z_s = ops.fine_z(z, w, int(N_importance), draw(), want_samples=True)
        raw_full, _ = manipulator_nerf(rays, pe, ve, model_fine, z_vals=z_full)
        _, _, _, acc = manipulator_render(raw_full, z_full, rays[1])
        return raw, z, rgb, z_s, acc

    ori_raw, ori_z, _, _, ori_acc = stage1(ori_rays)
    tar_raws, tar_zs, tar_samples, tar_accs = [], [], [], []
    tar_rgb = tar_acc = None
    for tar_rays in f_tar_rays:
        t_raw, t_z, tar_rgb, t_s, tar_acc = stage1(tar_rays)
        tar_raws.append(t_raw); tar_zs.append(t_z); tar_samples.append(t_s); tar_accs.append(tar_acc)

    ori_raw, _, _, _ = exchanger(ori_raw, tar_raws, ori_acc, tar_accs, labels, want_labels=False)

    # stage 2: re-sample on the edited coarse field, merge every sample set, query the fine field
    _, ori_w, _, _ = manipulator_render(ori_raw, ori_z, ori_rays[1])
    mids = .5 * (ori_z[..., 1:] + ori_z[..., :-1])
    ori_s = ops.sample_pdf(mids.contiguous(), ori_w[..., 1:-1].contiguous(), int(N_importance), draw())
    ori_z_all = ops.sort_cat([ori_z, ori_s] + tar_samples)
    ori_raw2 = None
    for idx, tar_rays in enumerate(f_tar_rays):
        if ori_raw2 is None:   # the reference recomputes this identical query once per target (:194)
            ori_raw2, _ = manipulator_nerf(ori_rays, pe, ve, model_fine, z_vals=ori_z_all)
        t_z_all = ops.sort_cat([tar_zs[idx], ori_s] + tar_samples)
        tar_raws[idx], _ = manipulator_nerf(tar_rays, pe, ve, model_fine, z_vals=t_z_all)
    ori_raw2, _, _, _ = exchanger(ori_raw2, tar_raws, ori_acc, tar_accs, labels, want_labels=False)
    final_rgb, _, _, final_ins = manipulator_render(ori_raw2, ori_z_all, ori_rays[1])
    return final_rgb, final_ins, tar_rgb, tar_acc
