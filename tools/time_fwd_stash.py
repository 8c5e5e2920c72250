"""Times the TRAINING forward (activation + mask stash) of the bf16 field query; argv: n_rays S K flags."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import _lib, ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
s = int(sys.argv[2]) if len(sys.argv) > 2 else 192
K = int(sys.argv[3]) if len(sys.argv) > 3 else 93
flag = int(sys.argv[4]) if len(sys.argv) > 4 else 0
ops.set_mode("bf16")
m = DM_NeRF(8, 256, 63, 27, [4], K); m.load_state_dict(orc.init_params(K, 1)); m = m.cuda()
o, d = torch.rand(n, 3, device="cuda"), torch.randn(n, 3, device="cuda")
z = torch.sort(torch.rand(n, s, device="cuda") * 11 + 4, -1)[0]
packed = m.packed(ops.get_mode())
_lib.lib().dmnerf_debug_flags(flag)
res = {}
for stash in (False, True):
    for _ in range(3): ops.mlp_fwd(o, d, z, packed, K, ops.get_mode(), stash=stash)
    ts = []
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ops.mlp_fwd(o, d, z, packed, K, ops.get_mode(), stash=stash); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    res[stash] = sorted(ts)[3]
pairs = n * s / 256
print(f"flags {flag} n={n} S={s} K={K}: inference {res[False]*1e3:.0f} us, training {res[True]*1e3:.0f} us "
      f"({res[True]/res[False]:.2f}x); rounds at 148 CTAs: {pairs/148:.2f}")
