"""Embedder / get_embedder / DM_NeRF with the reference's surface (networks/dm_nerf.py:8-106).

``DM_NeRF`` keeps the reference's module tree, so ``state_dict()`` keys and shapes are identical
(mlps.0..7, rgb_feature_linear, ins_feature_linear, rgb_feature_linears.0, ins_feature_linears.0,
density_linear, ins_linear, rgb_linear; dm_nerf.py:65-78) and reference checkpoints load.  The fp32
parameters are the master copy; packed K-major images for the kernels are derived caches refreshed
whenever a parameter's version counter moves (i.e. after every optimizer step).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops
from .._lib import MODE_BF16, MODE_F16, MODE_FP32, MODE_TC32


class Embedder:
    """[x, sin(2^k x), cos(2^k x)]_{k<L}; same kwargs as the reference (dm_nerf.py:8-38)."""

    def __init__(self, **kwargs):
        self.kwargs = kwargs
        d = kwargs.get("input_dims", 3)
        n = kwargs["num_freqs"]
        if (d != 3 or not kwargs.get("include_input", True) or not kwargs.get("log_sampling", True)
                or kwargs.get("max_freq_log2", n - 1) != n - 1):
            raise RuntimeError("Embedder: only the reference's configuration (3 inputs, include_input, "
                               "log-sampled bands 2^0..2^(L-1), sin/cos) is supported")
        self.num_freqs = n
        self.out_dim = d + 2 * d * n

    def embed(self, inputs: torch.Tensor) -> torch.Tensor:
        return ops.embed(inputs, self.num_freqs)


def get_embedder(multires, i=0):
    """dm_nerf.py:41-55."""
    if i == -1:
        return nn.Identity(), 3
    e = Embedder(include_input=True, input_dims=3, max_freq_log2=multires - 1, num_freqs=multires,
                 log_sampling=True, periodic_fns=[torch.sin, torch.cos])
    return e, e.out_dim


class DM_NeRF(nn.Module):
    def __init__(self, D=8, W=256, input_ch_pts=3, input_ch_views=3, skips=[4], ins_num=None):
        super().__init__()
        if (D, W, input_ch_pts, input_ch_views, list(skips)) != (8, 256, 63, 27, [4]) or ins_num is None:
            raise RuntimeError(
                "dm_nerf_b200.DM_NeRF supports the shape every shipped config builds: D=8, W=256, 63+27 input "
                f"channels, skips=[4] (got D={D}, W={W}, in={input_ch_pts}+{input_ch_views}, skips={skips}, "
                f"ins_num={ins_num}); there is no fallback path")
        self.skips = skips
        self.input_ch_pts = input_ch_pts
        self.input_ch_views = input_ch_views
        self.ins_num = int(ins_num)
        self.mlps = nn.ModuleList(
            [nn.Linear(input_ch_pts, W)]
            + [nn.Linear(W + input_ch_pts, W) if i in skips else nn.Linear(W, W) for i in range(D - 1)])
        self.rgb_feature_linear = nn.Linear(W, W)
        self.ins_feature_linear = nn.Linear(W, W)
        self.rgb_feature_linears = nn.ModuleList([nn.Linear(W + input_ch_views, W // 2)])
        self.ins_feature_linears = nn.ModuleList([nn.Linear(W, W // 2)])
        self.density_linear = nn.Linear(W, 1)
        self.ins_linear = nn.Linear(W // 2, ins_num + 1)
        self.rgb_linear = nn.Linear(W // 2, 3)
        self._packed = {}
        self._dirty = False

    # -- packed-weight cache ---------------------------------------------------------------------
    def _ordered_params(self):
        sd = dict(self.named_parameters())
        return [sd[k] for k in ops.PARAM_ORDER]

    def packed(self, mode: int) -> torch.Tensor:
        """Operand image of the current weights, cached on (data_ptr, version) of the 30 tensors.  A forward with
        gradients enabled marks the cache dirty: an optimizer step follows it, and not every optimizer moves the version
        counters (torch's fused multi-tensor Adam does not), so the next query always repacks."""
        if self._dirty:
            self._packed.clear()
            self._dirty = False
        params = self._ordered_params()
        key = tuple((p.data_ptr(), p._version) for p in params) + (ops.variant_selected(),)
        hit = self._packed.get(mode)
        if hit is None or hit[0] != key:
            hit = (key, ops.pack_weights(params, self.ins_num, mode))
            self._packed[mode] = hit
        return hit[1]

    def _wants_grad(self) -> bool:
        return torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())

    # -- queries ------------------------------------------------------------------------------------
    def query_rays(self, rays_o, rays_d, z_vals) -> torch.Tensor:
        """raw[N,S,5+K] for points rays_o + rays_d*z with both embeddings computed in-kernel
        (render.py:49-61 / manipulator.py:108-134 in one launch)."""
        mode = ops.get_mode()
        if self._wants_grad():
            packed = self.packed(mode)
            self._dirty = True
            return ops._MlpFn.apply(rays_o, rays_d, z_vals, None, self.ins_num, mode, packed, *self._ordered_params())
        return ops.mlp_fwd(rays_o, rays_d, z_vals, self.packed(mode), self.ins_num, mode)[0]

    def query_density(self, rays_o, rays_d, z_vals) -> torch.Tensor:
        """sigma[N,S] = ``query_rays(...)[..., 3]`` (same bits), inference only.  In the tensor-core modes the trunk and
        ``density_linear`` alone are evaluated (dm_nerf.py:82-92); the other modes take the column of the full query."""
        mode = ops.get_mode()
        if mode in (MODE_BF16, MODE_F16, MODE_TC32):
            return ops.mlp_density(rays_o, rays_d, z_vals, self.packed(mode), self.ins_num, mode)
        return ops.mlp_fwd(rays_o, rays_d, z_vals, self.packed(mode), self.ins_num, mode)[0][..., 3].contiguous()

    def forward(self, x):
        """x[..., 90] already embedded (dm_nerf.py:80-106)."""
        mode = ops.get_mode()
        if self._wants_grad():
            packed = self.packed(mode)
            self._dirty = True
            return ops._MlpFn.apply(None, None, None, x, self.ins_num, mode, packed, *self._ordered_params())
        return ops.mlp_fwd_embedded(x, self.packed(mode), self.ins_num, mode)[0]
