"""Rounding-faithful CPU model of the bf16 tensor-core MLP path -- TEST INFRASTRUCTURE ONLY.

Same network as ``dmnerf_oracle.mlp_forward`` (networks/dm_nerf.py:80-106) with bf16 rounding inserted exactly
where dm_nerf_b200/csrc/mlp_tc.cu and mlp_tc_bwd.cu round: weights (after the fp32 composition of the two
linear->linear pairs) and embeddings as MMA operands, every post-ReLU activation, and -- in the backward --
every pre-activation gradient G_l.  Accumulation stays fp32.  It exists because a ReLU unit whose
pre-activation is within bf16 noise of zero flips its mask between an fp32 and a bf16 forward; that moves a
parameter gradient by ~sqrt(flipped fraction) (several percent after 8 layers) without being an error of the
kernels.  Against THIS model the kernels must agree to accumulation-order level; against the fp32 oracle
only direction and scale are asserted.  Only tests/ may import it."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F
from torch import Tensor

import dmnerf_oracle as orc


def _bf(x: Tensor) -> Tensor:
    return x.bfloat16().float()


def rb(x: Tensor) -> Tensor:
    """bf16 rounding in the forward, identity in the backward."""
    return x + (_bf(x) - x).detach()


class _RoundGrad(torch.autograd.Function):
    """Identity in the forward; the gradient is rounded to bf16 (the stashed G_l operand)."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return _bf(g)


def mlp_forward_bf16(p: Dict[str, Tensor], x: Tensor) -> Tensor:
    pts, dirs = torch.split(x, [orc.POS_CH, orc.DIR_CH], dim=-1)
    pts, dirs = rb(pts), rb(dirs)
    h = pts
    pre = None
    for i in range(orc.TRUNK_DEPTH):
        pre = _RoundGrad.apply(F.linear(h, rb(p[f"mlps.{i}.weight"]), p[f"mlps.{i}.bias"]))
        h = rb(F.relu(pre))
        if i == orc.SKIP_AFTER:
            h = torch.cat([h, pts], -1)
    density = F.linear(F.relu(pre), p["density_linear.weight"], p["density_linear.bias"])   # fp32, un-rounded h7
    w1 = p["rgb_feature_linears.0.weight"]
    wc_rgb = rb(w1[:, :256] @ p["rgb_feature_linear.weight"])
    bc_rgb = w1[:, :256] @ p["rgb_feature_linear.bias"] + p["rgb_feature_linears.0.bias"]
    pre_rgb = F.linear(h, wc_rgb) + F.linear(dirs, rb(w1[:, 256:])) + bc_rgb
    rgb_hid = rb(F.relu(_RoundGrad.apply(pre_rgb)))
    w1i = p["ins_feature_linears.0.weight"]
    wc_ins = rb(w1i @ p["ins_feature_linear.weight"])
    bc_ins = w1i @ p["ins_feature_linear.bias"] + p["ins_feature_linears.0.bias"]
    ins_hid = rb(F.relu(_RoundGrad.apply(F.linear(h.detach(), wc_ins) + bc_ins)))           # dm_nerf.py:95
    rgb = F.linear(rgb_hid, rb(p["rgb_linear.weight"]), p["rgb_linear.bias"])
    ins = F.linear(ins_hid, rb(p["ins_linear.weight"]), p["ins_linear.bias"])
    return torch.cat([rgb, density, ins], -1)


def query_field_bf16(p: Dict[str, Tensor], rays_o: Tensor, rays_d: Tensor, z: Tensor) -> Tensor:
    viewdirs = rays_d / torch.norm(rays_d, dim=-1, keepdim=True)
    pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
    e_pos = orc.embed(torch.reshape(pts, [-1, 3]), 10)
    e_dir = orc.embed(torch.reshape(viewdirs[:, None].expand(pts.shape), [-1, 3]), 4)
    raw = mlp_forward_bf16(p, torch.cat([e_pos, e_dir], -1))
    return torch.reshape(raw, list(pts.shape[:-1]) + [raw.shape[-1]])
