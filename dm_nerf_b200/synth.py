"""Seeded synthetic inputs of the measured workloads: random-init weights, cameras, label maps.

Everything here is drawn from private CPU generators, so the same seed gives the same tensors on every machine
(the bench's GPU arm, its CPU arm and the parity tests all start from identical weights).  The product path does not
depend on ``oracle/``; ``tests/test_oracle_vs_golden.py`` pins ``init_state_dict`` to the oracle's initialiser.
"""
from __future__ import annotations

import math
from typing import Dict

import torch


def param_shapes(ins_num: int) -> Dict[str, tuple]:
    """state_dict keys/shapes of ``DM_NeRF(8, 256, 63, 27, [4], ins_num)`` in registration order (dm_nerf.py:65-78)."""
    out: Dict[str, tuple] = {}

    def lin(name, n_in, n_out):
        out[name + ".weight"] = (n_out, n_in)
        out[name + ".bias"] = (n_out,)
    lin("mlps.0", 63, 256)
    for i in range(1, 8):
        lin(f"mlps.{i}", 319 if i == 5 else 256, 256)
    lin("rgb_feature_linear", 256, 256)
    lin("ins_feature_linear", 256, 256)
    lin("rgb_feature_linears.0", 283, 128)
    lin("ins_feature_linears.0", 256, 128)
    lin("density_linear", 256, 1)
    lin("ins_linear", 128, ins_num + 1)
    lin("rgb_linear", 128, 3)
    return out


def init_state_dict(ins_num: int, seed: int, head_gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """``nn.Linear``'s default initialisation (U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias) from a private
    generator, in state_dict order.  ``head_gain`` scales the three output layers (the "peaky" variant of SURVEY 8(d):
    a trained network's logits are far from tied, a random one's are not)."""
    g = torch.Generator().manual_seed(seed)
    shapes = param_shapes(ins_num)
    out: Dict[str, torch.Tensor] = {}
    for name, shp in shapes.items():
        layer = name.rsplit(".", 1)[0]
        bound = 1.0 / math.sqrt(shapes[layer + ".weight"][1])
        t = (torch.rand(shp, generator=g) * 2.0 - 1.0) * bound
        if head_gain != 1.0 and layer in ("density_linear", "ins_linear", "rgb_linear"):
            t = t * head_gain
        out[name] = t
    return out


def pose_spherical(theta_deg: float, phi_deg: float, radius: float) -> torch.Tensor:
    """Camera-to-world of a camera on a sphere looking at the origin (the shape of datasets/loader_dmsr.py:29-34)."""
    t = torch.eye(4)
    t[2, 3] = radius
    ph, th = phi_deg / 180.0 * math.pi, theta_deg / 180.0 * math.pi
    rp = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0], [0, math.sin(ph), math.cos(ph), 0],
                       [0, 0, 0, 1]], dtype=torch.float32)
    rt = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0], [math.sin(th), 0, math.cos(th), 0],
                       [0, 0, 0, 1]], dtype=torch.float32)
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=torch.float32)
    return flip @ (rt @ rp @ t)


def dmsr_camera(H: int, W: int, view: int = 0):
    """DM-SR style intrinsics (loader_dmsr.py:171-172) and the ``view``-th pose of a 45-degree sweep at -65 degrees
    elevation, radius 7 (loader_dmsr.py:160-161)."""
    focal = 0.5 * W / math.tan(0.5 * 0.6911)
    K = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
    return K, pose_spherical(37.0 + 45.0 * view, -65.0, 7.0)


def scannet_camera(H: int = 968, W: int = 1296, view: int = 0):
    """ScanNet-shaped colour intrinsics (fx = fy = 1170, cx = 648, cy = 484) and a sweep pose."""
    K = torch.tensor([[1170.0, 0, 648.0], [0, 1170.0, 484.0], [0, 0, 1]], dtype=torch.float32)
    return K, pose_spherical(20.0 + 45.0 * view, -40.0, 4.0)


def block_labels(H: int, W: int, ins_num: int, device=None) -> torch.Tensor:
    """Instance mask of 8 x 8 blocks with labels in [0, ins_num)."""
    r = torch.arange(H, device=device)[:, None] // max(1, H // 8)
    c = torch.arange(W, device=device)[None] // max(1, W // 8)
    return (r * 8 + c) % ins_num
