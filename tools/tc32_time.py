"""Launch time of the DMNERF_MODE_TC32 forward (32768 rays x 192 samples, K=32) under debug flags given on the command
line (e.g. 0 64), each with the cycle counters of block 0.  python tools/tc32_time.py 0 64"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth, _lib
from dm_nerf_b200.networks.dm_nerf import DM_NeRF

dev, k, n, s = "cuda", 32, 32768, 192
m = DM_NeRF(8, 256, 63, 27, [4], k); m.load_state_dict(synth.init_state_dict(k, 11)); m = m.to(dev)
gen = torch.Generator().manual_seed(4)
o = ((torch.rand(n, 3, generator=gen) - 0.5) * 2).to(dev)
d = torch.randn(n, 3, generator=gen).to(dev)
z = torch.sort(torch.rand(n, s, generator=gen) * 11 + 4, -1)[0].to(dev)
ops.set_mode("tc32")
ref = None
for flag in [int(x) for x in sys.argv[1:]] or [0]:
    _lib.lib().dmnerf_debug_flags(flag | 16)
    with torch.no_grad():
        for _ in range(2): out = m.query_rays(o, d, z)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): m.query_rays(o, d, z)
        b.record(); torch.cuda.synchronize()
    if ref is None: ref = out
    buf = (C.c_longlong * 12)()
    _lib.lib().dmnerf_debug_counters_tc32(buf)
    tiles = (n * s / 128) / 148
    print(f"flags {flag}: {a.elapsed_time(b) / 5:.3f} ms  max|diff vs first| {float((out - ref).abs().max()):.2e}  per tile: issuer {buf[0] / tiles:.0f} "
          f"(w_full {buf[1] / tiles:.0f}, a_ready {buf[2] / tiles:.0f})  epi wait acc {buf[9] / tiles:.0f}  producer wait {buf[10] / tiles:.0f}", flush=True)
_lib.lib().dmnerf_debug_flags(0)
