"""The ray pipeline with the reference's surface (networks/render.py).

``dm_nerf`` (= north_star ``render_rays``) runs: optional stratified jitter -> coarse field query
(PE + MLP fused, one launch) -> compositing (one launch) -> coarse tail (pdf, inverse-CDF sampling,
sorted merge; one launch) -> fine field query -> compositing.  Five to six kernel launches per call
instead of the reference's several hundred ATen ops; no intermediate but ``raw``/``z`` (which the
callers consume) touches HBM.
"""
from __future__ import annotations

import torch

from .. import ops
from .._lib import COMP_DROP_EMPTY


def render_train(raw, z_vals, rays_d):
    """render.py:6-28 -> (rgb_map [N,3], weights [N,S], depth_map [N], ins_map [N,K]).
    Instance compositing uses detached weights (render.py:22-24) and drops the 'empty' class (:26)."""
    return ops.composite(raw, z_vals, rays_d, COMP_DROP_EMPTY, True)


raw2outputs = render_train


def _query(model, rays_o, rays_d, z_vals, position_embedder, view_embedder):
    if hasattr(model, "query_rays") and getattr(position_embedder, "num_freqs", None) == 10 \
            and getattr(view_embedder, "num_freqs", None) == 4:
        return model.query_rays(rays_o, rays_d, z_vals)
    raise RuntimeError("dm_nerf: models must be dm_nerf_b200 DM_NeRF modules with the L=10 / L=4 embedders "
                       "(get_embedder(10), get_embedder(4)); there is no generic fallback")


def dm_nerf(rays, position_embedder, view_embedder, model_coarse, model_fine, z_vals_coarse, args,
            t_rand=None, u=None, on_coarse=None, penalizer=None, on_fine=None):
    """render.py:31-96.  ``t_rand`` [N,Sc] / ``u`` [N,N_importance] inject the uniforms; when absent
    they are drawn with torch.rand in the reference's order (jitter first, then u).  ``on_coarse(ins_coarse)``, if
    given, is called as soon as the coarse pass is composited (the training step starts the coarse instance
    assignment on a side stream there, beside the fine pass); ``on_fine(ins_fine)`` likewise right after the fine
    compositing launch, before the fine penalizer's (grad mode may be off inside it).  ``penalizer=(tolerance, deta_w)`` evaluates
    ``ins_penalizer`` (penalizer.py:58-62) of both passes inside the compositing nodes and adds ``emptiness_coarse`` /
    ``emptiness_fine`` to the returned dict (same values as calling ``ins_penalizer`` on the outputs; one fused
    backward per pass instead of two gradients of ``raw`` that autograd then sums)."""
    def sliced(fn):
        if fn is None:
            return None
        return lambda ins: fn(ins[-args.N_ins:] if (args.is_train and args.N_ins is not None) else ins)

    def composite(raw, z, hook=None):
        if penalizer is not None and torch.is_grad_enabled() and raw.requires_grad:
            return ops.composite_penalized(raw, z, rays_d, COMP_DROP_EMPTY, True, penalizer[0], penalizer[1], hook)
        out = render_train(raw, z, rays_d)
        if hook is not None:
            hook(out[3])
        if penalizer is not None:
            return out + (ops.penalizer(raw, z, out[2].detach(), rays_d, penalizer[0], penalizer[1]),)
        return out + (None,)

    rays_o, rays_d = rays
    rays_o, rays_d = rays_o.float().contiguous(), rays_d.float().contiguous()
    z_vals_coarse = z_vals_coarse.float()
    n = rays_d.shape[0]
    if args.perturb > 0.:
        if t_rand is None:
            t_rand = torch.rand(z_vals_coarse.shape, device=rays_d.device)
        z_vals_coarse = ops.jitter_z(z_vals_coarse, t_rand)
    else:
        z_vals_coarse = z_vals_coarse.contiguous()
    raw_coarse = _query(model_coarse, rays_o, rays_d, z_vals_coarse, position_embedder, view_embedder)
    rgb_coarse, weights_coarse, depth_coarse, ins_coarse, pen_coarse = composite(raw_coarse, z_vals_coarse)
    if on_coarse is not None:
        on_coarse(ins_coarse[-args.N_ins:] if (args.is_train and args.N_ins is not None) else ins_coarse)
    det = (args.perturb == 0.)
    if det:
        u_arg = torch.linspace(0., 1., steps=args.N_importance, device=rays_d.device)
    else:
        u_arg = u if u is not None else torch.rand([n, args.N_importance], device=rays_d.device)
    z_vals_fine = ops.fine_z(z_vals_coarse, weights_coarse.detach(), int(args.N_importance), u_arg)
    raw_fine = _query(model_fine, rays_o, rays_d, z_vals_fine, position_embedder, view_embedder)
    rgb_fine, weights_fine, depth_fine, ins_fine, pen_fine = composite(raw_fine, z_vals_fine, sliced(on_fine))
    if args.is_train and args.N_ins is not None:
        ins_fine = ins_fine[-args.N_ins:]
        ins_coarse = ins_coarse[-args.N_ins:]
    extra = {} if penalizer is None else {'emptiness_coarse': pen_coarse, 'emptiness_fine': pen_fine}
    return {**extra, 'rgb_fine': rgb_fine, 'ins_fine': ins_fine, 'z_vals_fine': z_vals_fine, 'raw_fine': raw_fine,
            'raw_coarse': raw_coarse, 'rgb_coarse': rgb_coarse, 'ins_coarse': ins_coarse,
            'z_vals_coarse': z_vals_coarse, 'depth_fine': depth_fine, 'depth_coarse': depth_coarse}


render_rays = dm_nerf
