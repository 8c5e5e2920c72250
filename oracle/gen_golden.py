"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from
/root/reference, torch CPU fp32) on seeded inputs.  Run in the authoring container only:

    python oracle/gen_golden.py            # writes tests/golden/

The reference cannot travel to the GPU box, so the vectors are committed.  Weights are
not stored: they are regenerated from a seed by ``dmnerf_oracle.init_params`` (CPU
mt19937 generator -> machine independent) and loaded into the reference's ``DM_NeRF``
with ``load_state_dict``.  The reference's ``torch.rand`` draws (render.py:46,
helpers.py:135) are replaced by a queue of pre-drawn tensors that are stored in the
fixture, so every consumer sees identical uniforms.
"""
from __future__ import annotations

import os
import sys
import types
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("DMNERF_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, HERE)

import dmnerf_oracle as orc  # noqa: E402


def import_reference():
    """Put /root/reference on sys.path; stub the IO-only third-party modules that
    networks/manipulator.py imports at module top (manipulator.py:3-15) and that are
    absent here.  None of them is touched by the compute functions."""
    for name in ["imageio", "lpips", "skimage", "skimage.metrics", "open3d", "matplotlib",
                 "matplotlib.pyplot", "trimesh", "h5py", "configargparse"]:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if not hasattr(sys.modules["skimage"], "metrics"):
        sys.modules["skimage"].metrics = sys.modules["skimage.metrics"]
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, REF)
    import networks.dm_nerf as r_model
    import networks.render as r_render
    import networks.helpers as r_helpers
    import networks.penalizer as r_pen
    import networks.manipulator as r_mani
    torch.autograd.set_detect_anomaly(False)  # the reference switches it on at import (dm_nerf.py:5)
    return SimpleNamespace(model=r_model, render=r_render, helpers=r_helpers, pen=r_pen, mani=r_mani)


class RandQueue:
    """Stand-in for torch.rand inside the reference: hands out pre-drawn tensors in order."""

    def __init__(self):
        self.q = []
        self._orig = torch.rand

    def push(self, *ts):
        self.q.extend(ts)

    def __call__(self, *shape, **kw):
        if kw.get("generator") is not None or not self.q:
            return self._orig(*shape, **kw)
        t = self.q.pop(0)
        want = list(shape[0]) if len(shape) == 1 and not isinstance(shape[0], int) else list(shape)
        assert list(t.shape) == [int(s) for s in want], (t.shape, want)
        return t.clone()


def ref_models(R, ins_num, seed_c, seed_f, head_gain=1.0):
    pc = orc.init_params(ins_num, seed_c, head_gain)
    pf = orc.init_params(ins_num, seed_f, head_gain)
    mc = R.model.DM_NeRF(8, 256, 63, 27, [4], ins_num)
    mf = R.model.DM_NeRF(8, 256, 63, 27, [4], ins_num)
    mc.load_state_dict(pc)
    mf.load_state_dict(pf)
    return pc, pf, mc, mf


def synth_rays(n, seed, near_origin=True):
    g = torch.Generator().manual_seed(seed)
    o = (torch.rand(n, 3, generator=g) - 0.5) * 2.0
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True) * (0.8 + 0.6 * torch.rand(n, 1, generator=g))
    return torch.stack([o, d], 0)


ONLY = set(sys.argv[1:])          # `python oracle/gen_golden.py select_crop ...` rewrites just the named fixtures


def save(name, **arrs):
    if ONLY and name not in ONLY:
        return
    os.makedirs(OUT, exist_ok=True)
    conv = {}
    for k, v in arrs.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        conv[k] = np.asarray(v)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **conv)
    print("wrote", name, {k: tuple(v.shape) for k, v in conv.items()})


def main():
    torch.manual_seed(0)
    torch.set_num_threads(1)  # fixed reduction partitioning for the committed vectors
    R = import_reference()
    rq = RandQueue()
    torch.rand = rq

    # ---- a1 embedding -------------------------------------------------------------------
    pe, pe_dim = R.model.get_embedder(10, 0)
    ve, ve_dim = R.model.get_embedder(4, 0)
    assert (pe_dim, ve_dim) == (63, 27)
    g = torch.Generator().manual_seed(11)
    x = torch.cat([torch.tensor([[0.1, 0.2, 0.3], [15.0, -15.0, 7.5], [0.0, 1.0, -1.0]]),
                   (torch.rand(29, 3, generator=g) - 0.5) * 30.0], 0)
    save("embed", x=x, pos=pe.embed(x), dir=ve.embed(x))

    # ---- a2 MLP forward on embedded input ---------------------------------------------------
    for ins_num in (32, 16, 93):
        pc, _, mc, _ = ref_models(R, ins_num, 100 + ins_num, 200 + ins_num)
        g = torch.Generator().manual_seed(12)
        pts = (torch.rand(257, 3, generator=g) - 0.5) * 8.0
        dirs = torch.randn(257, 3, generator=g)
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        xin = torch.cat([pe.embed(pts), ve.embed(dirs)], -1)
        with torch.no_grad():
            out = mc(xin)
        save(f"mlp_k{ins_num}", pts=pts, dirs=dirs, out=out, seed=np.int64(100 + ins_num), ins_num=np.int64(ins_num))

    # ---- a5 / a8 compositing --------------------------------------------------------------
    g = torch.Generator().manual_seed(13)
    for tag, n, s, k in (("c64", 33, 64, 32), ("f192", 19, 192, 32), ("k93", 7, 192, 93), ("k16", 9, 320, 16)):
        raw = torch.randn(n, s, k + 5, generator=g) * 2.0
        raw[..., 3] = raw[..., 3] * 3.0
        z, _ = torch.sort(torch.rand(n, s, generator=g) * 11.0 + 4.0, -1)
        rd = torch.randn(n, 3, generator=g)
        a = R.render.render_train(raw, z, rd)
        b = R.mani.manipulator_render(raw, z, rd)
        save(f"composite_{tag}", raw=raw, z=z, rays_d=rd, rgb=a[0], weights=a[1], depth=a[2], ins=a[3],
             m_rgb=b[0], m_weights=b[1], m_depth=b[2], m_ins=b[3])
    # KATs
    n, s, k = 4, 8, 3
    raw = torch.zeros(n, s, k + 5)
    raw[1, 2, 3] = 1e6
    raw[1, 2, :3] = torch.tensor([0.3, -0.7, 1.1])
    raw[2, s - 1, 3] = 1e-3
    raw[2, s - 1, 4:] = torch.tensor([1.0, -2.0, 3.0, 0.5])
    raw[3, :, 3] = 0.7
    z = torch.linspace(2.0, 6.0, s).expand(n, s).contiguous()
    rd = torch.tensor([[0.0, 0.0, 1.0], [1.0, 2.0, 2.0], [0.0, 0.0, -2.0], [0.0, 0.0, 0.0]])
    a = R.render.render_train(raw, z, rd)
    save("composite_kat", raw=raw, z=z, rays_d=rd, rgb=a[0], weights=a[1], depth=a[2], ins=a[3])

    # ---- a4 sample_pdf --------------------------------------------------------------------
    g = torch.Generator().manual_seed(14)
    n = 41
    bins, _ = torch.sort(torch.rand(n, 63, generator=g) * 11.0 + 4.0, -1)
    w = torch.rand(n, 62, generator=g) ** 4
    w[0] = 1.0                      # uniform
    w[1] = 0.0                      # all zero -> 1e-5 floor
    w[2] = 0.0
    w[2, 17] = 5.0                  # single spike (denom < 1e-5 neighbours)
    w[3] = 1e-7
    u = torch.rand(n, 128, generator=g)
    u[4, 0] = 0.0
    u[4, 1] = 1.0 - 2.0 ** -24
    u[5, :] = torch.linspace(0, 1, 128)
    rq.push(u)
    zr = R.helpers.sample_pdf(bins, w, 128, det=False)
    zd = R.helpers.sample_pdf(bins, w, 128, det=True)
    _, inds_r, cdf = orc.sample_pdf(bins, w, 128, det=False, u=u, return_inds=True)
    _, inds_d, _ = orc.sample_pdf(bins, w, 128, det=True, return_inds=True)
    save("sample_pdf", bins=bins, weights=w, u=u, z_rand=zr, z_det=zd, inds_rand=inds_r, inds_det=inds_d, cdf=cdf)

    # ---- a6 dm_nerf end to end -------------------------------------------------------------
    for tag, ins_num, n, perturb, gain in (("det_k32", 32, 17, 0.0, 8.0), ("rand_k32", 32, 13, 1.0, 8.0),
                                           ("det_k16", 16, 9, 0.0, 1.0)):
        pc, pf, mc, mf = ref_models(R, ins_num, 300 + ins_num, 400 + ins_num, gain)
        rays = synth_rays(n, 21 + n)
        zc = R.helpers.z_val_sample(n, 4.0, 15.0, 64)
        args = SimpleNamespace(perturb=perturb, N_importance=128, is_train=False, N_ins=None)
        g = torch.Generator().manual_seed(15)
        t_rand = torch.rand(n, 64, generator=g)
        uu = torch.rand(n, 128, generator=g)
        if perturb > 0:
            rq.push(t_rand, uu)
        with torch.no_grad():
            info = R.render.dm_nerf(rays, pe, ve, mc, mf, zc, args)
        assert not rq.q
        save(f"dm_nerf_{tag}", rays=rays, t_rand=t_rand, u=uu, perturb=np.float32(perturb),
             ins_num=np.int64(ins_num), seed_c=np.int64(300 + ins_num), seed_f=np.int64(400 + ins_num),
             head_gain=np.float32(gain), near=np.float32(4.0), far=np.float32(15.0),
             **{k: v for k, v in info.items()})

    # ---- a7 penalizer (+ gradient) -----------------------------------------------------------
    g = torch.Generator().manual_seed(16)
    n, s, k = 23, 64, 32
    raw = (torch.randn(n, s, k + 5, generator=g) * 2.0).requires_grad_(True)
    z = R.helpers.z_val_sample(n, 4.0, 15.0, s).contiguous()
    depth = torch.rand(n, generator=g) * 11.0 + 4.0
    depth[0] = 3.0       # in front of the first sample: "before" mask empty for that ray
    depth[1] = 16.0      # behind the last sample
    rd = torch.randn(n, 3, generator=g)
    args = SimpleNamespace(tolerance=0.05, deta_w=0.05)
    loss = R.pen.ins_penalizer(raw, z, depth, rd, args)
    loss.backward()
    save("penalizer", raw=raw.detach(), z=z, depth=depth, rays_d=rd, loss=loss.detach(), d_raw=raw.grad,
         tolerance=np.float32(0.05), deta_w=np.float32(0.05))
    # all masks empty on the "before" side: every depth in front of the samples
    raw2 = (torch.randn(5, s, k + 5, generator=g)).requires_grad_(True)
    depth2 = torch.full((5,), 1.0)
    loss2 = R.pen.ins_penalizer(raw2, z[:5], depth2, rd[:5], args)
    loss2.backward()
    save("penalizer_empty", raw=raw2.detach(), z=z[:5], depth=depth2, rays_d=rd[:5], loss=loss2.detach(),
         d_raw=raw2.grad, tolerance=np.float32(0.05), deta_w=np.float32(0.05))

    # ---- a10 exchanger --------------------------------------------------------------------------
    g = torch.Generator().manual_seed(17)
    n, s, k = 37, 64, 32
    ori = torch.randn(n, s, k + 5, generator=g) * 6.0
    tars = [torch.randn(n, s, k + 5, generator=g) * 6.0 for _ in range(2)]
    # make labels 5 and 9 frequent, and plant saturated logits (sigmoid -> 1.0f ties)
    for t in [ori] + tars:
        pick = torch.rand(n, s, generator=g)
        t[..., 4 + 5] += (pick < 0.25) * 12.0
        t[..., 4 + 9] += ((pick > 0.25) & (pick < 0.45)) * 12.0
        t[0, :8, 4 + 20] = 40.0
        t[0, :8, 4 + 7] = 30.0
    ori_acc = torch.sigmoid(torch.randn(n, k + 1, generator=g) * 3.0)
    tar_accs = [torch.sigmoid(torch.randn(n, k + 1, generator=g) * 3.0) for _ in range(2)]
    for a in [ori_acc] + tar_accs:
        a[: n // 2, 5] = 0.999
        a[n // 2:, 9] = 0.998
    ori_in = ori.clone()
    tars_in = [t.clone() for t in tars]
    o_raw, _, o_lab, t_lab = R.mani.exchanger(ori, tars, ori_acc, tar_accs, [5, 9])
    save("exchanger", ori_raw=ori_in, tar_raw0=tars_in[0], tar_raw1=tars_in[1], ori_acc=ori_acc,
         tar_acc0=tar_accs[0], tar_acc1=tar_accs[1], move_labels=np.array([5, 9]),
         out_raw=o_raw, out_label=o_lab, tar_label=t_lab)

    # ---- a11 manipulator end to end ----------------------------------------------------------------
    ins_num, n = 32, 9
    pc, pf, mc, mf = ref_models(R, ins_num, 500, 600, 8.0)
    ori_rays = synth_rays(n, 31)
    tar_rays = synth_rays(n, 32)
    tar_rays = (ori_rays + 0.05 * tar_rays)[None]            # [T=1, 2, n, 3], a nearby camera
    g = torch.Generator().manual_seed(18)
    us = [torch.rand(n, 128, generator=g) for _ in range(3)]
    rq.push(*[u.clone() for u in us])
    args = SimpleNamespace(N_samples=64, N_importance=128, near=4.0, far=15.0, target_label=[8])
    with torch.no_grad():
        f_rgb, f_ins, t_rgb, t_acc = R.mani.manipulator(pe, ve, mc, mf, ori_rays, tar_rays, args)
    assert not rq.q
    save("manipulator", ori_rays=ori_rays, tar_rays=tar_rays, u0=us[0], u1=us[1], u2=us[2],
         final_rgb=f_rgb, final_ins=f_ins, tar_rgb=t_rgb, tar_acc=t_acc, target_label=np.array([8]),
         seed_c=np.int64(500), seed_f=np.int64(600), head_gain=np.float32(8.0), ins_num=np.int64(32))

    # ---- a12 training-step gradients -----------------------------------------------------------------
    ins_num, n = 32, 24
    pc, pf, mc, mf = ref_models(R, ins_num, 700, 800, 1.0)
    rays = synth_rays(n, 41)
    zc = R.helpers.z_val_sample(n, 4.0, 15.0, 64)
    g = torch.Generator().manual_seed(19)
    t_rand = torch.rand(n, 64, generator=g)
    uu = torch.rand(n, 128, generator=g)
    tgt_rgb = torch.rand(n, 3, generator=g)
    tgt_ins = torch.rand(n, ins_num, generator=g)
    rq.push(t_rand, uu)
    args = SimpleNamespace(perturb=1.0, N_importance=128, is_train=True, N_ins=None, tolerance=0.05, deta_w=0.05)
    info = R.render.dm_nerf(rays, pe, ve, mc, mf, zc, args)
    loss = ((info["rgb_fine"] - tgt_rgb) ** 2).mean() + ((info["rgb_coarse"] - tgt_rgb) ** 2).mean()
    loss = loss + ((info["ins_fine"] - tgt_ins) ** 2).mean() + ((info["ins_coarse"] - tgt_ins) ** 2).mean()
    loss = loss + R.pen.ins_penalizer(info["raw_coarse"], info["z_vals_coarse"], info["depth_coarse"], rays[1], args)
    loss = loss + R.pen.ins_penalizer(info["raw_fine"], info["z_vals_fine"], info["depth_fine"], rays[1], args)
    loss.backward()
    grads = {}
    for tag, m in (("c", mc), ("f", mf)):
        for name, prm in m.named_parameters():
            gflat = prm.grad.reshape(-1)
            grads[f"{tag}.{name}.norm"] = gflat.double().norm().float()
            grads[f"{tag}.{name}.sub"] = gflat[::97].clone()
    save("train_grads", rays=rays, t_rand=t_rand, u=uu, tgt_rgb=tgt_rgb, tgt_ins=tgt_ins, loss=loss.detach(),
         seed_c=np.int64(700), seed_f=np.int64(800), ins_num=np.int64(32), **grads)

    # ---- adjacent: get_rays_k -----------------------------------------------------------------------
    H, W = 6, 8
    focal = 0.5 * W / np.tan(0.5 * 0.6911)
    K = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
    c2w = orc.pose_spherical(37.0, -65.0, 7.0)
    ro, rd = R.helpers.get_rays_k(H, W, K, c2w)
    save("rays_k", K=K, c2w=c2w, rays_o=ro, rays_d=rd, H=np.int64(H), W=np.int64(W))

    # ---- adjacent (8f-2): Hungarian-matched instance loss (+ gradient) ----------------------------------
    import networks.evaluator as r_eval
    g = torch.Generator().manual_seed(21)
    for tag, kk, nn, n_lab in (("k32", 32, 257, 9), ("k93", 93, 130, 93)):
        pred = torch.sigmoid(torch.randn(nn, kk, generator=g) * 2.0).requires_grad_(True)
        lab = torch.randint(0, n_lab, (nn,), generator=g)
        out = r_eval.ins_criterion(pred, lab, kk)
        out[0].backward()
        save(f"ins_criterion_{tag}", pred=pred.detach(), labels=lab, ins_num=np.int64(kk),
             losses=torch.stack([o.reshape(()).float() for o in out]).detach(), grad=pred.grad)

    # ---- adjacent (8f-3): pixel selection of the training loops (helpers.py:64-111), numpy draws seeded ----------
    H, W = 12, 16
    focal = 0.5 * W / np.tan(0.5 * 0.6911)
    K = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
    c2w = orc.pose_spherical(111.0, -65.0, 7.0)
    g = torch.Generator().manual_seed(31)
    rgb = torch.rand(H, W, 3, generator=g)
    ins_t = torch.randint(0, 7, (H, W), generator=g)
    crop = torch.zeros(H, W); crop[2:10, 3:14] = 1
    ins_index = np.where((crop.reshape(-1).numpy() == 1) & (ins_t.reshape(-1).numpy() < 4))[0]
    np.random.seed(1234)
    tc, ti, br, n_ins = R.helpers.get_select_crop(rgb, c2w, K, ins_t, ins_index, crop, 40)
    np.random.seed(1234)
    fc, fi, fr = R.helpers.get_select_full(rgb, c2w, K, ins_t, 40)
    save("select_crop", K=K, c2w=c2w, rgb=rgb, ins_target=ins_t, crop_mask=crop, ins_index=ins_index, N_train=np.int64(40),
         seed=np.int64(1234), target_c=tc, target_i=ti, rays=br, N_ins=np.int64(n_ins), full_c=fc, full_i=fi, full_rays=fr)

    # ---- a10 on realistic logits: the coarse field's own raw and the fine pass's accumulated maps, as manipulator.py
    # :141-177 hands them to exchanger (no planted ties; the move label is the most frequent pixel label) -------------
    ins_num, n = 32, 96
    pc, pf, mc, mf = ref_models(R, ins_num, 510, 610, 8.0)
    ori_rays = synth_rays(n, 51)
    tar_rays = ori_rays + 0.05 * synth_rays(n, 52)
    g = torch.Generator().manual_seed(23)
    with torch.no_grad():
        def stage(rays, u):
            raw, z = R.mani.manipulator_nerf(rays, pe, ve, mc, 64, 4.0, 15.0)
            _, w, _, _ = R.mani.manipulator_render(raw, z, rays[1])
            rq.push(u)
            zs = R.helpers.sample_pdf(.5 * (z[..., 1:] + z[..., :-1]), w[..., 1:-1], 128).detach()
            zf, _ = torch.sort(torch.cat([z, zs], -1), -1)
            raw_f, _ = R.mani.manipulator_nerf(rays, pe, ve, mf, z_vals=zf)
            _, _, _, acc = R.mani.manipulator_render(raw_f, zf, rays[1])
            return raw, acc
        o_raw_in, o_acc = stage(ori_rays, torch.rand(n, 128, generator=g))
        t_raw_in, t_acc = stage(tar_rays, torch.rand(n, 128, generator=g))
        move = int(torch.bincount(torch.argmax(torch.sigmoid(o_acc[..., :-1]), -1), minlength=ins_num).argmax())
        o_raw, _, o_lab, t_lab = R.mani.exchanger(o_raw_in.clone(), [t_raw_in.clone()], o_acc, [t_acc], [move])
    save("exchanger_real", ori_raw=o_raw_in, tar_raw0=t_raw_in, ori_acc=o_acc, tar_acc0=t_acc,
         move_labels=np.array([move]), out_raw=o_raw, out_label=o_lab, tar_label=t_lab)

    # ---- a11 on a batch large enough for a statistic (the 9-ray fixture above stays as the smoke case) -----------------
    ins_num, n = 32, 96
    pc, pf, mc, mf = ref_models(R, ins_num, 500, 600, 8.0)
    ori_rays = synth_rays(n, 61)
    tar_rays = (ori_rays + 0.05 * synth_rays(n, 62))[None]
    g = torch.Generator().manual_seed(28)
    us = [torch.rand(n, 128, generator=g) for _ in range(3)]
    rq.push(*[u.clone() for u in us])
    with torch.no_grad():
        stage1_raw, _ = R.mani.manipulator_nerf(ori_rays, pe, ve, mc, 64, 4.0, 15.0)
        _, _, _, acc1 = R.mani.manipulator_render(stage1_raw, R.mani.manipulator_nerf(ori_rays, pe, ve, mc, 64, 4.0, 15.0)[1],
                                                  ori_rays[1])
        move = int(torch.bincount(torch.argmax(acc1[..., :-1], -1), minlength=ins_num).argmax())
        args = SimpleNamespace(N_samples=64, N_importance=128, near=4.0, far=15.0, target_label=[move])
        f_rgb, f_ins, t_rgb, t_acc = R.mani.manipulator(pe, ve, mc, mf, ori_rays, tar_rays, args)
    assert not rq.q
    save("manipulator_n96", ori_rays=ori_rays, tar_rays=tar_rays, u0=us[0], u1=us[1], u2=us[2],
         final_rgb=f_rgb, final_ins=f_ins, tar_rgb=t_rgb, tar_acc=t_acc, target_label=np.array([move]),
         seed_c=np.int64(500), seed_f=np.int64(600), head_gain=np.float32(8.0), ins_num=np.int64(32))

    torch.rand = rq._orig


if __name__ == "__main__":
    # `gen_golden.py only <prefix>`: write just the fixtures whose name starts with <prefix> (adding one must not
    # rewrite the committed others)
    if len(sys.argv) > 2 and sys.argv[1] == "only":
        _save, _prefix = save, sys.argv[2]

        def save(name, **arrs):  # noqa: F811
            if name.startswith(_prefix):
                _save(name, **arrs)
    main()
