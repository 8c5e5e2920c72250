"""CPU oracle for DM-NeRF's per-ray hot path -- TEST INFRASTRUCTURE, NOT PRODUCT.

A functional, torch-CPU-fp32 restatement of what the reference computes on the path
SURVEY.md section 8(a) names.  The reference's arithmetic for this path *is* PyTorch
ATen, so the restatement uses the same ATen primitives in the same order and is
bit-identical to the reference on the same torch build (pinned by
``tests/test_oracle_vs_golden.py`` against fixtures made by ``oracle/gen_golden.py``
from the imported reference, torch 2.11.0 CPU).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The shipped package
(``dm_nerf_b200``) never does; it fails loudly when its CUDA library is missing.

Every function cites the reference lines (relative to /root/reference) it follows.
Randomness is always *injected* (``t_rand``, ``u``) so that parity runs share uniforms.
"""
from __future__ import annotations

import math

import numpy as np
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# ----------------------------------------------------------------------------------------
# a1  positional embedding                                   networks/dm_nerf.py:8-55
# ----------------------------------------------------------------------------------------


def embed(x: Tensor, n_freqs: int) -> Tensor:
    """[x, sin(x*2^0), cos(x*2^0), ..., sin(x*2^(L-1)), cos(x*2^(L-1))] along the last dim.

    Frequencies are 2**linspace(0, L-1, L) (dm_nerf.py:24-25, log_sampling=True), the
    input comes first (dm_nerf.py:17-19) and each band contributes sin then cos
    (dm_nerf.py:29-32, periodic_fns order at :51).
    """
    bands = 2.0 ** torch.linspace(0.0, float(n_freqs - 1), steps=n_freqs)
    parts = [x]
    for f in bands:
        parts.append(torch.sin(x * f))
        parts.append(torch.cos(x * f))
    return torch.cat(parts, -1)


# ----------------------------------------------------------------------------------------
# a2  DM_NeRF MLP                                            networks/dm_nerf.py:58-106
# ----------------------------------------------------------------------------------------

TRUNK_DEPTH = 8
WIDTH = 256
SKIP_AFTER = 4          # cat([h, pts]) happens after layer index 4 (dm_nerf.py:86-87)
POS_CH = 63
DIR_CH = 27


def param_shapes(ins_num: int) -> Dict[str, Tuple[int, ...]]:
    """state_dict layout of DM_NeRF(8, 256, 63, 27, [4], ins_num) (dm_nerf.py:65-78)."""
    shp: Dict[str, Tuple[int, ...]] = {}
    for i in range(TRUNK_DEPTH):
        if i == 0:
            k = POS_CH
        elif i == SKIP_AFTER + 1:
            k = WIDTH + POS_CH
        else:
            k = WIDTH
        shp[f"mlps.{i}.weight"] = (WIDTH, k)
        shp[f"mlps.{i}.bias"] = (WIDTH,)
    shp["rgb_feature_linear.weight"] = (WIDTH, WIDTH)
    shp["rgb_feature_linear.bias"] = (WIDTH,)
    shp["ins_feature_linear.weight"] = (WIDTH, WIDTH)
    shp["ins_feature_linear.bias"] = (WIDTH,)
    shp["rgb_feature_linears.0.weight"] = (WIDTH // 2, WIDTH + DIR_CH)
    shp["rgb_feature_linears.0.bias"] = (WIDTH // 2,)
    shp["ins_feature_linears.0.weight"] = (WIDTH // 2, WIDTH)
    shp["ins_feature_linears.0.bias"] = (WIDTH // 2,)
    shp["density_linear.weight"] = (1, WIDTH)
    shp["density_linear.bias"] = (1,)
    shp["ins_linear.weight"] = (ins_num + 1, WIDTH // 2)
    shp["ins_linear.bias"] = (ins_num + 1,)
    shp["rgb_linear.weight"] = (3, WIDTH // 2)
    shp["rgb_linear.bias"] = (3,)
    return shp


def init_params(ins_num: int, seed: int, head_gain: float = 1.0) -> Dict[str, Tensor]:
    """Random-init weights with nn.Linear's default scheme (kaiming-uniform(a=sqrt 5) ==
    U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias), drawn from a private
    generator so fixtures do not depend on module construction order.  ``head_gain``
    scales the three output layers ("peaky" variant of SURVEY 8(d)) so argmax is not
    near-tied."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, Tensor] = {}
    for name, shp in param_shapes(ins_num).items():
        layer = name.rsplit(".", 1)[0]
        fan_in = param_shapes(ins_num)[layer + ".weight"][1]
        bound = 1.0 / math.sqrt(fan_in)
        t = (torch.rand(shp, generator=g) * 2.0 - 1.0) * bound
        if head_gain != 1.0 and layer in ("density_linear", "ins_linear", "rgb_linear"):
            t = t * head_gain
        out[name] = t
    return out


def mlp_forward(p: Dict[str, Tensor], x: Tensor) -> Tensor:
    """x[M, 90] (63 pos-embedding | 27 dir-embedding) -> raw[M, 5+ins_num] = cat(rgb3, sigma1, ins).

    dm_nerf.py:80-106.  ReLU precedes the skip concat, so layer 5 sees [relu(h) | pts]
    (:85-87); rgb branch: linear (no act) -> cat dirs -> 283->128 ReLU -> 3 (:89-93,102);
    instance branch reads h.detach() (:95-99,103); density straight off h (:101).
    """
    pts, dirs = torch.split(x, [POS_CH, DIR_CH], dim=-1)
    h = pts
    for i in range(TRUNK_DEPTH):
        h = F.relu(F.linear(h, p[f"mlps.{i}.weight"], p[f"mlps.{i}.bias"]))
        if i == SKIP_AFTER:
            h = torch.cat([h, pts], -1)
    rf = F.linear(h, p["rgb_feature_linear.weight"], p["rgb_feature_linear.bias"])
    rf = torch.cat([rf, dirs], -1)
    rf = F.relu(F.linear(rf, p["rgb_feature_linears.0.weight"], p["rgb_feature_linears.0.bias"]))
    jf = F.linear(h.detach(), p["ins_feature_linear.weight"], p["ins_feature_linear.bias"])
    jf = F.relu(F.linear(jf, p["ins_feature_linears.0.weight"], p["ins_feature_linears.0.bias"]))
    sigma = F.linear(h, p["density_linear.weight"], p["density_linear.bias"])
    rgb = F.linear(rf, p["rgb_linear.weight"], p["rgb_linear.bias"])
    ins = F.linear(jf, p["ins_linear.weight"], p["ins_linear.bias"])
    return torch.cat([rgb, sigma, ins], -1)


def query_field(p: Dict[str, Tensor], rays_o: Tensor, rays_d: Tensor, z: Tensor,
                l_pos: int = 10, l_dir: int = 4) -> Tensor:
    """Points o + d*z, both embeddings, one MLP call -> raw[N, S, C].
    render.py:37-38,49-61 (identical again at :71-83) and manipulator.py:108-134."""
    viewdirs = rays_d / torch.norm(rays_d, dim=-1, keepdim=True)
    viewdirs = torch.reshape(viewdirs, [-1, 3]).float()
    pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
    flat = torch.reshape(pts, [-1, 3])
    e_pos = embed(flat, l_pos)
    d_flat = torch.reshape(viewdirs[:, None].expand(pts.shape), [-1, 3])
    e_dir = embed(d_flat, l_dir)
    raw = mlp_forward(p, torch.cat([e_pos, e_dir], -1))
    return torch.reshape(raw, list(pts.shape[:-1]) + [raw.shape[-1]])


# ----------------------------------------------------------------------------------------
# a3  coarse depths                                          networks/helpers.py:114-119
# ----------------------------------------------------------------------------------------


def z_val_sample(n_rays: int, near: float, far: float, n_samples: int) -> Tensor:
    """near + linspace(0,1,S)*(far-near), one row expanded to every ray (helpers.py:114-119)."""
    lo = near * torch.ones(size=(n_rays, 1))
    hi = far * torch.ones(size=(n_rays, 1))
    t = torch.linspace(0.0, 1.0, steps=n_samples)
    return (lo + t * (hi - lo)).expand([n_rays, n_samples])


def z_val_lerp(n_rays: int, near: float, far: float, n_samples: int) -> Tensor:
    """The manipulator's algebraically different form near*(1-t)+far*t (manipulator.py:116-120)."""
    lo = near * torch.ones(size=(n_rays, 1))
    hi = far * torch.ones(size=(n_rays, 1))
    t = torch.linspace(0.0, 1.0, steps=n_samples)
    return (lo * (1.0 - t) + hi * t).expand([n_rays, n_samples])


# ----------------------------------------------------------------------------------------
# a4  hierarchical sampler                                   networks/helpers.py:123-155
# ----------------------------------------------------------------------------------------


def sample_pdf(bins: Tensor, weights: Tensor, n_samples: int, det: bool = False,
               u: Optional[Tensor] = None, return_inds: bool = False):
    """Inverse-CDF sampling.  ``u`` [N, n_samples] replaces the reference's torch.rand
    draw (helpers.py:135) when det is False.

    w += 1e-5; pdf = w / sum(w); cdf = [0, cumsum(pdf)] (:125-128); u = linspace(0,1,n)
    when det (:131-133); inds = searchsorted(cdf, u, right=True) (:139); below =
    max(inds-1, 0), above = min(inds, len(cdf)-1) (:140-141); denom < 1e-5 -> 1 (:150-151);
    z = bins_b + (u - cdf_b)/denom * (bins_a - bins_b) (:152-153).
    """
    w = weights + 1e-5
    pdf = w / torch.sum(w, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    if det:
        u = torch.linspace(0.0, 1.0, steps=n_samples).expand(list(cdf.shape[:-1]) + [n_samples])
    else:
        assert u is not None, "oracle.sample_pdf: pass the uniforms explicitly when det=False"
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b = torch.gather(cdf, -1, below)
    cdf_a = torch.gather(cdf, -1, above)
    bin_b = torch.gather(bins, -1, below)
    bin_a = torch.gather(bins, -1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    z = bin_b + t * (bin_a - bin_b)
    if return_inds:
        return z, inds, cdf
    return z


# ----------------------------------------------------------------------------------------
# a5 / a8  volume compositing                 networks/render.py:6-28, manipulator.py:86-105
# ----------------------------------------------------------------------------------------


def composite(raw: Tensor, z: Tensor, rays_d: Tensor, drop_empty: bool, detach_ins: bool):
    """alpha = 1-exp(-relu(sigma)*delta), T = exclusive cumprod(1-alpha+1e-10), w = alpha*T,
    rgb = sum w*sigmoid(raw_rgb), depth = sum w*z, ins = sigmoid(sum w*raw_ins).

    ``drop_empty`` / ``detach_ins`` True  -> render_train      (render.py:6-28)
    both False                            -> manipulator_render (manipulator.py:86-105)
    The last interval is 1e10 (render.py:10) and every interval is scaled by |rays_d| (:12).
    """
    step = z[..., 1:] - z[..., :-1]
    step = torch.cat([step, torch.Tensor([1e10]).expand(step[..., :1].shape)], -1)
    step = step * torch.norm(rays_d[..., None, :], dim=-1)
    colour = torch.sigmoid(raw[..., :3])
    logits = raw[..., 4:]
    alpha = 1.0 - torch.exp(-F.relu(raw[..., 3]) * step)
    trans = torch.cumprod(torch.cat([torch.ones((alpha.shape[0], 1)), 1.0 - alpha + 1e-10], -1), -1)[:, :-1]
    w = alpha * trans
    rgb_map = torch.sum(w[..., None] * colour, -2)
    depth_map = torch.sum(w * z, -1)
    w_ins = w.clone().detach() if detach_ins else w
    ins_map = torch.sigmoid(torch.sum(w_ins[..., None] * logits, -2))
    if drop_empty:
        ins_map = ins_map[..., :-1]
    return rgb_map, w, depth_map, ins_map


def render_train(raw, z, rays_d):
    return composite(raw, z, rays_d, drop_empty=True, detach_ins=True)


def manipulator_render(raw, z, rays_d):
    return composite(raw, z, rays_d, drop_empty=False, detach_ins=False)


# ----------------------------------------------------------------------------------------
# a6  the ray pipeline                                       networks/render.py:31-96
# ----------------------------------------------------------------------------------------


def dm_nerf(rays: Tensor, p_coarse, p_fine, z_coarse: Tensor, *, perturb: float, n_importance: int,
            t_rand: Optional[Tensor] = None, u: Optional[Tensor] = None,
            is_train: bool = False, n_ins: Optional[int] = None) -> Dict[str, Tensor]:
    """render.py:31-96 with the two torch.rand draws (:46 and helpers.py:135) injected."""
    rays_o, rays_d = rays
    if perturb > 0.0:
        mids = 0.5 * (z_coarse[..., 1:] + z_coarse[..., :-1])
        upper = torch.cat([mids, z_coarse[..., -1:]], -1)
        lower = torch.cat([z_coarse[..., :1], mids], -1)
        assert t_rand is not None
        z_coarse = lower + (upper - lower) * t_rand
    raw_c = query_field(p_coarse, rays_o, rays_d, z_coarse)
    rgb_c, w_c, depth_c, ins_c = render_train(raw_c, z_coarse, rays_d)
    z_mid = 0.5 * (z_coarse[..., 1:] + z_coarse[..., :-1])
    z_new = sample_pdf(z_mid, w_c[..., 1:-1], n_importance, det=(perturb == 0.0), u=u).detach()
    z_fine, _ = torch.sort(torch.cat([z_coarse, z_new], -1), -1)
    raw_f = query_field(p_fine, rays_o, rays_d, z_fine)
    rgb_f, w_f, depth_f, ins_f = render_train(raw_f, z_fine, rays_d)
    if is_train and n_ins is not None:
        ins_f = ins_f[-n_ins:]
        ins_c = ins_c[-n_ins:]
    return {"rgb_fine": rgb_f, "ins_fine": ins_f, "z_vals_fine": z_fine, "raw_fine": raw_f,
            "raw_coarse": raw_c, "rgb_coarse": rgb_c, "ins_coarse": ins_c, "z_vals_coarse": z_coarse,
            "depth_fine": depth_f, "depth_coarse": depth_c, "weights_coarse": w_c, "weights_fine": w_f}


# ----------------------------------------------------------------------------------------
# a7  emptiness penalizer                                    networks/penalizer.py:5-62
# ----------------------------------------------------------------------------------------


def emptiness_penalizer(raw: Tensor, z: Tensor, depth: Tensor, rays_d: Tensor,
                        tolerance: float, deta_w: float) -> Tensor:
    """penalizer.py:5-55.  ``depth`` is [N, 1] (ins_penalizer unsqueezes and detaches it, :59)."""
    sd_h = torch.Tensor([0.4])
    sd_w = torch.Tensor([deta_w])
    length = torch.norm(rays_d[..., None, :], dim=-1)
    front = (depth - tolerance) * length
    back = (depth + tolerance) * length
    surf = depth * length
    along = z * length
    gap = surf - along
    bell = torch.exp(-(gap ** 2) / (2 * (sd_w ** 2))) / (sd_h * torch.sqrt(torch.Tensor([2 * math.pi]))) + 1e-8
    air = 1 - bell
    m_before = (along < front).type(torch.float32)
    m_after = (along > back).type(torch.float32)
    m_middle = 1 - (m_after + m_before)
    prob = torch.sigmoid(raw[..., 4:])
    onehot = torch.zeros_like(prob)
    onehot[..., -1] = 1
    bce = -onehot * torch.log(prob + 1e-8) - (1 - onehot) * torch.log(1 - prob + 1e-8)
    bce = bce * (air * m_before)[..., None]
    loss_before = torch.sum(bce) / (prob.shape[-1] * torch.maximum(torch.sum(m_before), torch.tensor([1e-8])))
    p_empty = prob[..., -1]
    zero_t = torch.zeros_like(p_empty)
    mid = -zero_t * torch.log(p_empty + 1e-8) - (1 - zero_t) * torch.log(1 - p_empty + 1e-8)
    mid = mid * (bell * m_middle)
    loss_middle = torch.sum(mid) / torch.maximum(torch.sum(m_middle), torch.tensor([1e-8]))
    return loss_before + loss_middle


def ins_penalizer(raw, z, depth, rays_d, tolerance: float, deta_w: float) -> Tensor:
    """penalizer.py:58-62."""
    return emptiness_penalizer(raw, z, depth[..., None].detach(), rays_d, tolerance, deta_w)


# ----------------------------------------------------------------------------------------
# a10  exchanger                                             networks/manipulator.py:18-83
# ----------------------------------------------------------------------------------------


def exchanger(ori_raw: Tensor, tar_raws: List[Tensor], ori_acc: Tensor, tar_accs: List[Tensor],
              move_labels: Sequence[int]):
    """Per-sample label logic; mutates ``ori_raw`` in place like the reference.

    labels = argmax(sigmoid(logits)) (first max wins); the per-ray label comes from the
    already-sigmoided accumulated map without its last channel, sigmoided again
    (manipulator.py:19-26).  Per move label: relabel occluded samples (:31-36), fillings
    (:40-42), same relabel on the target (:44-60), op code from [tar==m] + 2*[ori==m]
    (:64-74), then copy / zero in that order (:77-81).
    """
    lab = torch.argmax(torch.sigmoid(ori_raw[..., 4:]), dim=-1)
    pix = torch.argmax(torch.sigmoid(ori_acc[..., :-1]), dim=-1)
    pix = pix[:, None].repeat(1, lab.shape[-1])
    tar_lab_out = None
    for k, m in enumerate(move_labels):
        t_raw = tar_raws[k]
        t_acc = tar_accs[k]
        hide = (pix != m) * (lab == m)
        lab[hide == True] = pix[hide == True]  # noqa: E712
        fill = (pix == m) * (lab != m)
        t_lab = torch.argmax(torch.sigmoid(t_raw[..., 4:]), dim=-1)
        tar_lab_out = t_lab
        t_pix = torch.argmax(torch.sigmoid(t_acc[..., :-1]), dim=-1)
        t_pix = t_pix[:, None].repeat(1, t_lab.shape[-1])
        t_hide = (t_pix != m) * (t_lab == m)
        t_lab[t_hide == True] = t_pix[t_hide == True]  # noqa: E712
        code = (t_lab == m).long() + 2 * (lab == m).long()
        swap = (code == 1) | (code == 3)
        kill = code == 2
        ori_raw[fill] = t_raw[fill]
        ori_raw[swap] = t_raw[swap]
        ori_raw[kill] = ori_raw[kill] * 0
    return ori_raw, tar_raws, lab, tar_lab_out


# ----------------------------------------------------------------------------------------
# a9 / a11  manipulator                                      networks/manipulator.py:108-205
# ----------------------------------------------------------------------------------------


def manipulator(p_coarse, p_fine, ori_rays: Tensor, tar_rays_list: Tensor, *, n_samples: int,
                n_importance: int, near: float, far: float, target_label: Sequence[int],
                us: Sequence[Tensor]):
    """manipulator.py:137-205.  ``us`` are the uniforms in the order the reference draws
    them: ori stage-1 (:148), one per target (:170), ori stage-2 (:187)."""
    us = list(us)
    n_rays = ori_rays[1].shape[0]

    def coarse(rays):
        zc = z_val_lerp(n_rays, near, far, n_samples)
        return query_field(p_coarse, rays[0], rays[1], zc), zc

    def refine(rays, zc, w, u):
        mid = 0.5 * (zc[..., 1:] + zc[..., :-1])
        zs = sample_pdf(mid, w[..., 1:-1], n_importance, det=False, u=u)
        zf, _ = torch.sort(torch.cat([zc, zs], dim=-1), dim=-1)
        return zs, zf

    ori_raw, ori_z = coarse(ori_rays)
    _, ori_w, _, _ = manipulator_render(ori_raw, ori_z, ori_rays[1])
    _, ori_zf = refine(ori_rays, ori_z, ori_w, us.pop(0))
    ori_raw_f = query_field(p_fine, ori_rays[0], ori_rays[1], ori_zf)
    _, _, _, ori_acc = manipulator_render(ori_raw_f, ori_zf, ori_rays[1])

    tar_raws, tar_zs, tar_samples, tar_accs = [], [], [], []
    tar_rgb = tar_acc = None
    for rays in tar_rays_list:
        t_raw, t_z = coarse(rays)
        tar_raws.append(t_raw)
        tar_zs.append(t_z)
        tar_rgb, t_w, _, _ = manipulator_render(t_raw, t_z, rays[1])
        t_s, t_zf = refine(rays, t_z, t_w, us.pop(0))
        t_raw_f = query_field(p_fine, rays[0], rays[1], t_zf)
        _, _, _, tar_acc = manipulator_render(t_raw_f, t_zf, rays[1])
        tar_samples.append(t_s)
        tar_accs.append(tar_acc)

    ori_raw, _, _, _ = exchanger(ori_raw, tar_raws, ori_acc, tar_accs, target_label)

    _, ori_w, _, _ = manipulator_render(ori_raw, ori_z, ori_rays[1])
    mid = 0.5 * (ori_z[..., 1:] + ori_z[..., :-1])
    ori_s = sample_pdf(mid, ori_w[..., 1:-1], n_importance, det=False, u=us.pop(0))
    all_tar_s = torch.cat(tar_samples, dim=-1)
    ori_z, _ = torch.sort(torch.cat([ori_z, ori_s, all_tar_s], dim=-1), dim=-1)
    for k, rays in enumerate(tar_rays_list):
        ori_raw = query_field(p_fine, ori_rays[0], ori_rays[1], ori_z)
        t_z, _ = torch.sort(torch.cat([tar_zs[k], ori_s, all_tar_s], dim=-1), dim=-1)
        tar_raws[k] = query_field(p_fine, rays[0], rays[1], t_z)
    ori_raw, _, _, _ = exchanger(ori_raw, tar_raws, ori_acc, tar_accs, target_label)
    final_rgb, _, _, final_ins = manipulator_render(ori_raw, ori_z, ori_rays[1])
    return final_rgb, final_ins, tar_rgb, tar_acc


# ----------------------------------------------------------------------------------------
# adjacent: pinhole rays                                     networks/helpers.py:50-61
# ----------------------------------------------------------------------------------------


def get_rays_k(H: int, W: int, K: Tensor, c2w: Tensor):
    """dirs = [(i-cx)/fx, (j-cy)/fy, K22]; rays_d = sum(dirs[..., None, :] * R, -1);
    rays_o = t broadcast (helpers.py:50-61)."""
    i, j = torch.meshgrid(torch.linspace(0, W - 1, W), torch.linspace(0, H - 1, H), indexing="ij")
    i = i.t()
    j = j.t()
    dirs = torch.stack([(i - K[0, 2]) / K[0, 0], (j - K[1, 2]) / K[1, 1], K[2, 2] * torch.ones_like(i)], -1)
    rays_d = torch.sum(dirs[..., None, :] * c2w[:3, :3], -1)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    return rays_o, rays_d


def get_select_full(rgb: Tensor, pose: Tensor, K: Tensor, ins_target: Tensor, n_train: int):
    """helpers.py:99-111: n_train distinct pixels (numpy draw), their rays / colours / labels."""
    H, W, _ = rgb.shape
    rays_o, rays_d = get_rays_k(H, W, K, pose)
    idx = torch.from_numpy(np.random.choice(H * W, size=[n_train], replace=False))
    return (rgb.reshape(-1, 3)[idx], ins_target.reshape(-1)[idx],
            torch.stack([rays_o.reshape(-1, 3)[idx], rays_d.reshape(-1, 3)[idx]], 0))


def get_select_crop(rgb: Tensor, pose: Tensor, K: Tensor, ins_target: Tensor, ins_index, crop_mask: Tensor, n_train: int):
    """helpers.py:64-96: 30 % labelled pixels (drawn from ins_index) placed LAST, the rest from the crop mask.  The
    unlabelled draw indexes ``crop_indices`` with positions drawn over the SIZE of the set difference (:84-85)."""
    n_ins = min(int(n_train * 0.3), len(ins_index))
    n_rgb = n_train - n_ins
    H, W, _ = rgb.shape
    rays_o, rays_d = get_rays_k(H, W, K, pose)
    crop_indices = np.where(crop_mask.reshape(-1).numpy() == 1)[0]
    ins_index = np.asarray(ins_index)
    labeled = ins_index[np.random.choice(ins_index.shape[0], size=[n_ins], replace=False)]
    n_unl = len(set(crop_indices.tolist()) - set(labeled.tolist()))
    unlabeled = crop_indices[np.random.choice(n_unl, size=[n_rgb], replace=False)]
    idx = torch.from_numpy(np.concatenate([unlabeled, labeled]).astype(np.int64))
    return (rgb.reshape(-1, 3)[idx], ins_target.reshape(-1)[idx[n_rgb:]],
            torch.stack([rays_o.reshape(-1, 3)[idx], rays_d.reshape(-1, 3)[idx]], 0), n_ins)


# ----------------------------------------------------------------------------------------
# Losses around the path that the timed train step includes (stock torch; evaluator.py)
# ----------------------------------------------------------------------------------------


def img2mse(x, y):
    """evaluator.py:11."""
    return torch.mean((x - y) ** 2)


def pose_spherical(theta_deg: float, phi_deg: float, radius: float) -> Tensor:
    """Synthetic camera used by bench/tests (datasets/loader_dmsr.py:29-34 shape of pose)."""
    t = torch.eye(4)
    t[2, 3] = radius
    ph = phi_deg / 180.0 * math.pi
    rp = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0],
                       [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1]], dtype=torch.float32)
    th = theta_deg / 180.0 * math.pi
    rt = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0],
                       [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1]], dtype=torch.float32)
    c2w = rt @ rp @ t
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=torch.float32)
    return flip @ c2w
