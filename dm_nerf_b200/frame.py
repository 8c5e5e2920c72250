"""Whole-frame driver: the chunk loop of render_test (networks/tester.py:54-84) with
rays generated on the device per chunk, outputs written into preallocated frame buffers (no
torch.cat growth) and an optional uint8 label map.  Rays shard over ranks by contiguous pixel
blocks (parallel.py); this function renders pixels [pix0, pix0+count) of one H x W frame."""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Optional

import torch

from . import ops
from .networks.render import dm_nerf


def render_pixels(position_embedder, view_embedder, model_coarse, model_fine, hwk, c2w, near: float, far: float,
                  n_samples: int = 64, n_importance: int = 128, pix0: int = 0, count: Optional[int] = None,
                  chunk: int = 32768, rays: Optional[torch.Tensor] = None, labels: bool = True,
                  out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """Deterministic (perturb=0) render of a pixel range.  Returns rgb [n,3], depth [n], ins [n,K]
    and (optionally) labels uint8 [n].  ``rays`` [2,n,3] overrides in-kernel ray generation."""
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
    with torch.no_grad():
        for a in range(0, count, chunk):
            b = min(count, a + chunk)
            if rays is not None:
                ro, rd = rays[0, a:b], rays[1, a:b]
            else:
                ro, rd = ops.rays_k(int(H), int(W), K, c2w, pix0 + a, b - a, dev)
            info = dm_nerf((ro, rd), position_embedder, view_embedder, model_coarse, model_fine,
                           z_row.expand(b - a, n_samples), args)
            out["rgb"][a:b] = info["rgb_fine"]
            out["depth"][a:b] = info["depth_fine"]
            out["ins"][a:b] = info["ins_fine"]
            if labels:
                out["labels"][a:b] = ops.label_argmax(info["ins_fine"])
    return out
