"""Stage-by-stage error of the fp32 CUDA path against the CPU oracle for a dm_nerf call (diagnostic)."""
import os
import sys
from types import SimpleNamespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc  # noqa: E402
from dm_nerf_b200 import ops  # noqa: E402
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder  # noqa: E402
from dm_nerf_b200.networks.render import dm_nerf, render_train  # noqa: E402
from dm_nerf_b200.networks.helpers import z_val_sample  # noqa: E402


def main():
    k, n = 32, 48
    gain = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
    normalize = len(sys.argv) > 2 and sys.argv[2] == "norm"
    pc, pf = orc.init_params(k, 1, gain), orc.init_params(k, 2, gain)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(pc); mf.load_state_dict(pf)
    mc, mf = mc.cuda(), mf.cuda()
    g = torch.Generator().manual_seed(0)
    o = (torch.rand(n, 3, generator=g) - 0.5) * 2
    d = torch.randn(n, 3, generator=g)
    if normalize:
        d = d / d.norm(dim=-1, keepdim=True)
    rays = torch.stack([o, d], 0)
    args = SimpleNamespace(perturb=0.0, N_importance=128, is_train=False, N_ins=None)
    pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
    with torch.no_grad():
        ref = orc.dm_nerf(rays, pc, pf, orc.z_val_sample(n, 4.0, 15.0, 64), perturb=0.0, n_importance=128)
        out = dm_nerf(rays.cuda(), pe, ve, mc, mf, z_val_sample(n, 4.0, 15.0, 64), args)
        for key in sorted(ref):
            if key in out:
                e = (out[key].cpu() - ref[key]).abs()
                print(f"{key:14s} max {float(e.max()):.3e} mean {float(e.mean()):.3e} ref-absmax {float(ref[key].abs().max()):.3e}")
        # fine stage on the ORACLE's z: isolates MLP + compositing from the sampler
        zf = ref["z_vals_fine"].cuda()
        raw_f = mf.query_rays(o.cuda(), d.cuda(), zf)
        e = (raw_f.cpu() - ref["raw_fine"]).abs()
        print(f"raw_fine@ref-z  max {float(e.max()):.3e} mean {float(e.mean()):.3e}; sigma ch max {float(e[..., 3].max()):.3e}")
        rgb, w, dep, ins = render_train(raw_f, zf, d.cuda())
        print("rgb_fine@ref-z  max", float((rgb.cpu() - ref["rgb_fine"]).abs().max()),
              "depth", float((dep.cpu() - ref["depth_fine"]).abs().max()))
        # compositing alone on the oracle's raw
        rgb2, w2, dep2, ins2 = render_train(ref["raw_fine"].cuda(), zf, d.cuda())
        print("composite(ref raw) rgb max", float((rgb2.cpu() - ref["rgb_fine"]).abs().max()),
              "depth", float((dep2.cpu() - ref["depth_fine"]).abs().max()))
        # the point positions
        pts_ref = o[:, None, :] + d[:, None, :] * ref["z_vals_fine"][:, :, None]
        print("max |pts|", float(pts_ref.abs().max()), " |d| max", float(d.norm(dim=-1).max()))


if __name__ == "__main__":
    main()
