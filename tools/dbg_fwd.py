import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import _lib, ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF
ops.set_mode("bf16")
m = DM_NeRF(8, 256, 63, 27, [4], 32); m.load_state_dict(orc.init_params(32, 1)); m = m.cuda()
n, s = 32768, 192
o, d = torch.rand(n, 3, device="cuda"), torch.randn(n, 3, device="cuda")
z = torch.sort(torch.rand(n, s, device="cuda") * 11 + 4, -1)[0]
with torch.no_grad():
    for _ in range(2):
        m.query_rays(o, d, z)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); m.query_rays(o, d, z); b.record(); torch.cuda.synchronize()
print("ms", a.elapsed_time(b))
buf = (C.c_longlong * 8)()
_lib.lib().dmnerf_debug_counters(buf)
names = ["mma total", "mma wait w_full", "mma wait h_ready", "epi total", "epi wait acc_done", "producer wait w_empty"]
pairs = (n * s / 256) / 148
for nm, v in zip(names, buf):
    print(f"{nm:24s} {v:12d}  per pair {v / pairs:10.0f}")
