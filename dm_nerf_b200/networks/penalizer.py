"""Emptiness penalizer with the reference's surface (networks/penalizer.py)."""
from __future__ import annotations

from .. import ops


def emptiness_penalizer(raw, z_vals, depths, rays_d, tolerance, deta_w):
    """penalizer.py:5-55; depths is [N,1] (or [N])."""
    return ops.penalizer(raw, z_vals, depths, rays_d, tolerance, deta_w)


def ins_penalizer(raw, z_vals, depth, rays_d, args):
    """penalizer.py:58-62 (depth is detached there)."""
    return ops.penalizer(raw, z_vals, depth.detach(), rays_d, args.tolerance, args.deta_w)
