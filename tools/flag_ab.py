"""A/B of forward-kernel debug flags on 32768 rays x S (K=32): bit-equality and launch time.
python tools/flag_ab.py [mode] [S] [flags ...]   e.g.  fp16 192 0 512"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth, _lib
from dm_nerf_b200.networks.dm_nerf import DM_NeRF
mode = sys.argv[1] if len(sys.argv) > 1 else "fp16"
s = int(sys.argv[2]) if len(sys.argv) > 2 else 192
flags = [int(x) for x in sys.argv[3:]] or [0, 512]
dev, k, n = "cuda", 32, 30720
m = DM_NeRF(8, 256, 63, 27, [4], k); m.load_state_dict(synth.init_state_dict(k, 11)); m = m.to(dev)
gen = torch.Generator().manual_seed(4)
ops.set_mode(mode)
for nn, ss in ((3, 77), (1000, 64), (700, 192)):
    o = ((torch.rand(nn, 3, generator=gen) - 0.5) * 2).to(dev); d = torch.randn(nn, 3, generator=gen).to(dev)
    z = torch.sort(torch.rand(nn, ss, generator=gen) * 11 + 4, -1)[0].to(dev)
    outs = []
    for f in flags:
        _lib.lib().dmnerf_debug_flags(f)
        with torch.no_grad(): outs.append(m.query_rays(o, d, z).clone())
    print(f"n={nn} s={ss}: equal={all(torch.equal(outs[0], x) for x in outs[1:])}", flush=True)
o = ((torch.rand(n, 3, generator=gen) - 0.5) * 2).to(dev); d = torch.randn(n, 3, generator=gen).to(dev)
z = torch.sort(torch.rand(n, s, generator=gen) * 11 + 4, -1)[0].to(dev)
ref = None
for f in flags * 3:
    _lib.lib().dmnerf_debug_flags(f | 16)
    with torch.no_grad():
        for _ in range(2): out = m.query_rays(o, d, z)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): m.query_rays(o, d, z)
        b.record(); torch.cuda.synchronize()
    if ref is None: ref = out
    buf = (C.c_longlong * 8)()
    _lib.lib().dmnerf_debug_counters(buf)
    pairs = (n * s / 256) / 148
    print(f"flags {f}: {a.elapsed_time(b) / 5:.3f} ms equal={torch.equal(out, ref)} per pair: " + " ".join(f"{v / pairs:.0f}" for v in buf[:6]), flush=True)
_lib.lib().dmnerf_debug_flags(0)
