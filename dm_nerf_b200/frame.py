"""Whole-frame driver: the chunk loop of render_test (networks/tester.py:54-84) with
rays generated on the device per chunk, outputs written into preallocated frame buffers (no
torch.cat growth) and an optional uint8 label map.  Rays shard over ranks by contiguous pixel
blocks (parallel.py); this function renders pixels [pix0, pix0+count) of one H x W frame."""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Optional

import torch

import numpy as np

from . import ops
from .networks.manipulator import manipulator
from .networks.render import dm_nerf, render_train


def render_pixels(position_embedder, view_embedder, model_coarse, model_fine, hwk, c2w, near: float, far: float,
                  n_samples: int = 64, n_importance: int = 128, pix0: int = 0, count: Optional[int] = None,
                  chunk: int = 32768, rays: Optional[torch.Tensor] = None, labels: bool = True,
                  out: Optional[Dict[str, torch.Tensor]] = None, balance: bool = True,
                  lean_coarse: bool = True, paced: bool = True) -> Dict[str, torch.Tensor]:
    """Deterministic (perturb=0) render of a pixel range.  Returns rgb [n,3], depth [n], ins [n,K]
    and (optionally) labels uint8 [n].  ``rays`` [2,n,3] overrides in-kernel ray generation.
    ``chunk`` bounds the rays per pass (raw_fine scratch = chunk * 192 * C * 4 bytes); with ``balance`` the range is cut
    into equal passes of at most ``chunk`` rays instead of full chunks plus a short tail (the persistent MLP kernels
    deal 256-point tile pairs round-robin to the SMs: a 5,632-ray tail costs whole rounds for a fraction of the work).
    ``paced`` (with ``lean_coarse``, in-kernel ray generation and ``balance``): the whole range goes through ONE
    ``dmnerf_render_full`` call -- same kernels, same results bit for bit, no interpreter or staging copy between launches."""
    H, W, K = hwk
    dev = next(model_fine.parameters()).device
    if count is None:
        count = (rays.shape[1] if rays is not None else int(H) * int(W) - pix0)
    kins = model_fine.ins_num
    if out is None:
        out = {"rgb": torch.empty((count, 3), device=dev), "depth": torch.empty((count,), device=dev),
               "ins": torch.empty((count, kins), device=dev)}
        if labels:
            out["labels"] = torch.empty((count,), dtype=torch.uint8, device=dev)
    args = SimpleNamespace(perturb=0.0, N_importance=n_importance, is_train=False, N_ins=None)
    t = torch.linspace(0., 1., steps=n_samples, device=dev)
    z_row = (near + t * (far - near))[None]
    if balance and count > 0:
        passes = -(-count // chunk)
        chunk = -(-count // passes)
        chunk = min(count, (chunk + 127) // 128 * 128)          # whole 128-point tiles per pass
    # lean_coarse: evaluate only the density trunk in the coarse pass (see below); needs our DM_NeRF modules
    lean = bool(lean_coarse) and hasattr(model_coarse, "query_density") and hasattr(model_fine, "query_rays") \
        and getattr(position_embedder, "num_freqs", None) == 10 and getattr(view_embedder, "num_freqs", None) == 4
    u_det = torch.linspace(0., 1., steps=n_importance, device=dev)
    if lean and rays is None and paced and balance and count > 0 and model_coarse.ins_num == kins:
        # every launch of every pass enqueued by ONE C call, compositing straight into the frame buffers
        mode = ops.get_mode()
        with torch.no_grad():
            ops.render_full(int(H), int(W), K, c2w, pix0, count, z_row, int(n_importance), model_coarse.packed(mode),
                            model_fine.packed(mode), kins, mode, chunk, out, labels, u_row=u_det)
        return out
    with torch.no_grad():
        for a in range(0, count, chunk):
            b = min(count, a + chunk)
            if rays is not None:
                ro, rd = rays[0, a:b], rays[1, a:b]
            else:
                ro, rd = ops.rays_k(int(H), int(W), K, c2w, pix0 + a, b - a, dev)
            if lean:
                # what tester.py:73-76 keeps is the FINE pass; of the coarse pass only the weights matter (they place the
                # fine samples, render.py:63-70): density-only trunk -> weights -> fine depths, then the fine pass as in
                # dm_nerf.  Same fine outputs bit for bit (tested), ~17 % off the coarse launch, no raw_coarse in HBM.
                rd_c, ro_c = rd.float().contiguous(), ro.float().contiguous()
                z_c = z_row.expand(b - a, n_samples).contiguous()
                sigma = model_coarse.query_density(ro_c, rd_c, z_c)
                w_c = ops.weights_from_sigma(sigma, z_c, rd_c, model_coarse.ins_num + 5)
                z_f = ops.fine_z(z_c, w_c, int(n_importance), u_det)
                raw_f = model_fine.query_rays(ro_c, rd_c, z_f)
                rgb_f, _, depth_f, ins_f = render_train(raw_f, z_f, rd_c)
            else:
                info = dm_nerf((ro, rd), position_embedder, view_embedder, model_coarse, model_fine,
                               z_row.expand(b - a, n_samples), args)
                rgb_f, depth_f, ins_f = info["rgb_fine"], info["depth_fine"], info["ins_fine"]
            out["rgb"][a:b] = rgb_f
            out["depth"][a:b] = depth_f
            out["ins"][a:b] = ins_f
            if labels:
                out["labels"][a:b] = ops.label_argmax(ins_f)
    return out


def deform_offsets(H: int, W: int, deform_func: str, amplitude: float = 0.0) -> torch.Tensor:
    """The per-pixel x-offset of the target rays' origins in ``mani_mode == 'deform'`` (manipulator.py:396-429): a
    function of the image row only, repeated along the row.  ``amplitude`` is the view's ``deform_v[i]`` ('sin' only).
    Returns float64 [H*W] like the reference (added to the fp32 origins)."""
    v = np.linspace(1, H, H)
    if deform_func == "sin":
        v = np.sin(((8 * np.pi) / 400) * v) * amplitude
    elif deform_func == "ex":
        v = np.exp(-1 * v / 50)
    elif deform_func == "linear":
        v = (v - 200) / 215
    elif deform_func == "abs_linear":
        v = np.abs(v - 200) / 200
    elif deform_func == "ln":
        v = np.log(v / 200)
    else:
        raise ValueError(f"unknown deform_func {deform_func!r}")
    return torch.from_numpy(np.repeat(v[:, None], W, axis=-1).reshape(-1))


def manipulate_pixels(position_embedder, view_embedder, model_coarse, model_fine, hwk, ori_pose, targets, args,
                      pix0: int = 0, count: Optional[int] = None, chunk: int = 16384,
                      out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """One edited view: the chunk loop of ``manipulator_demo`` (manipulator.py:383-469) over pixels [pix0, pix0+count).
    ``targets``: one dict per moved object with ``tar_id`` and either ``transformation`` (4x4; the target rays come
    from ``transformation @ ori_pose``, :431-433) or ``deform_func`` (+ ``amplitude``; origins shifted along x, :396-429).
    Rays are generated on the device per chunk and results land in preallocated buffers (no torch.cat growth).
    Returns rgb [n,3], ins [n,K+1], tar_rgb [n,3], tar_ins [n,K+1] and labels uint8 [n] (argmax over all K+1 channels of
    ins, K = "empty", as the demo's label image at manipulator.py:478)."""
    H, W, K = hwk
    dev = next(model_fine.parameters()).device
    if count is None:
        count = int(H) * int(W) - pix0
    kins = model_fine.ins_num
    if out is None:
        out = {"rgb": torch.empty((count, 3), device=dev), "ins": torch.empty((count, kins + 1), device=dev),
               "tar_rgb": torch.empty((count, 3), device=dev), "tar_ins": torch.empty((count, kins + 1), device=dev),
               "labels": torch.empty((count,), dtype=torch.uint8, device=dev)}
    ori_pose = torch.as_tensor(ori_pose, dtype=torch.float32)
    margs = SimpleNamespace(N_samples=args.N_samples, N_importance=args.N_importance, near=args.near, far=args.far,
                            target_label=[int(t["tar_id"]) for t in targets])
    tar_poses, offsets = [], []
    for t in targets:
        if t.get("deform_func") is not None:
            tar_poses.append(ori_pose)
            offsets.append(deform_offsets(int(H), int(W), t["deform_func"], float(t.get("amplitude", 0.0))).to(dev))
        else:
            tar_poses.append(torch.as_tensor(t["transformation"], dtype=torch.float32).cpu() @ ori_pose.cpu())
            offsets.append(None)
    with torch.no_grad():
        for a in range(0, count, chunk):
            b = min(count, a + chunk)
            ro, rd = ops.rays_k(int(H), int(W), K, ori_pose, pix0 + a, b - a, dev)
            tars = []
            for pose, off in zip(tar_poses, offsets):
                if off is None:
                    to, td = ops.rays_k(int(H), int(W), K, pose, pix0 + a, b - a, dev)
                else:
                    to, td = ro.clone(), rd
                    to[:, 0] = (to[:, 0] + off[pix0 + a:pix0 + b]).float()          # fp32 + fp64 -> fp32, as :427
                tars.append(torch.stack([to, td], 0))
            rgb, ins, tar_rgb, tar_ins = manipulator(position_embedder, view_embedder, model_coarse, model_fine,
                                                     torch.stack([ro, rd], 0), torch.stack(tars, 0), margs)
            out["rgb"][a:b], out["ins"][a:b], out["tar_rgb"][a:b], out["tar_ins"][a:b] = rgb, ins, tar_rgb, tar_ins
            out["labels"][a:b] = ops.label_argmax(ins)       # torch.argmax(ins, -1) over all K+1 channels (:478)
    return out


def query_occupancy(model_fine, points: torch.Tensor, voxel_size: float, chunk: int = 1 << 22) -> torch.Tensor:
    """The field query of ``mesh_main`` (tools/mesh_generator.py:36-62): ``occ = 1 - exp(-relu(sigma) * voxel_size)`` at
    arbitrary points (there: a 256^3 grid, embedded with zero view directions and sent through ``model_fine`` in N_test
    chunks that are ``torch.cat``-ed).  Density does not depend on the view direction (dm_nerf.py:88-92 takes it from
    the trunk, before the direction enters), so the points go through the fused ray kernel as one-sample rays
    ``o = point, d = e_z, z = 0`` (``o + d*0`` is exact): no [M,90] embedding in HBM, results in one preallocated buffer."""
    pts = points.reshape(-1, 3).float().contiguous()
    m = pts.shape[0]
    occ = torch.empty((m,), dtype=torch.float32, device=pts.device)
    ez = torch.tensor([0.0, 0.0, 1.0], device=pts.device)
    with torch.no_grad():
        for a in range(0, m, chunk):
            b = min(m, a + chunk)
            raw = model_fine.query_rays(pts[a:b], ez.expand(b - a, 3).contiguous(),
                                        torch.zeros((b - a, 1), dtype=torch.float32, device=pts.device))
            occ[a:b] = 1.0 - torch.exp(-torch.relu(raw[:, 0, 3]) * voxel_size)
    return occ.reshape(points.shape[:-1])
