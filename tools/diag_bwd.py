"""Per-parameter gradient error of the bf16 tensor-core backward vs the CPU oracle (diagnostic)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
import dmnerf_oracle_bf16 as orb
from dm_nerf_b200 import ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF

k, n, s = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
p = {kk: v.requires_grad_(True) for kk, v in orc.init_params(k, 9).items()}
m = DM_NeRF(8, 256, 63, 27, [4], k); m.load_state_dict(orc.init_params(k, 9)); m = m.cuda()
gen = torch.Generator().manual_seed(5)
o = (torch.rand(n, 3, generator=gen) - 0.5) * 2
d = torch.randn(n, 3, generator=gen)
z, _ = torch.sort(torch.rand(n, s, generator=gen) * 11 + 4, -1)
g_out = torch.randn(n, s, k + 5, generator=gen)
emu = len(sys.argv) > 4
ref = (orb.query_field_bf16 if emu else orc.query_field)(p, o, d, z)
(ref * g_out).sum().backward()
ops.set_mode("bf16")
got = m.query_rays(o.cuda(), d.cuda(), z.cuda())
(got * g_out.cuda()).sum().backward()
torch.cuda.synchronize()
print("raw err", float((got.detach().cpu() - ref.detach()).abs().max()))
for name, prm in m.named_parameters():
    a, b = prm.grad.detach().cpu().double().reshape(-1), p[name].grad.double().reshape(-1)
    rel = float((a - b).norm() / b.norm().clamp_min(1e-30))
    cos = float((a @ b) / (a.norm() * b.norm()).clamp_min(1e-30))
    print(f"{name:32s} rel {rel:.4f} cos {cos:.6f} |ref| {float(b.norm()):.3e} |got| {float(a.norm()):.3e}")
