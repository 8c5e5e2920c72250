"""Time the bf16 field query (a1+a2+a9) on 32768 rays x 192 samples; argv[1] = debug flags (0 lockstep, 4 pipelined, 8 pair)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import _lib, ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF
flag = int(sys.argv[1]) if len(sys.argv) > 1 else 0
K = int(sys.argv[2]) if len(sys.argv) > 2 else 32
ops.set_mode("bf16")
m = DM_NeRF(8, 256, 63, 27, [4], K); m.load_state_dict(orc.init_params(K, 1)); m = m.cuda()
n, s = 32768, 192
o, d = torch.rand(n, 3, device="cuda"), torch.randn(n, 3, device="cuda")
z = torch.sort(torch.rand(n, s, device="cuda") * 11 + 4, -1)[0]
with torch.no_grad():
    ref = m.query_rays(o, d, z)
    _lib.lib().dmnerf_debug_flags(flag)
    for _ in range(2):
        out = m.query_rays(o, d, z)
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); m.query_rays(o, d, z); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
print(f"flags {flag} K {K}: ms {min(ts):.3f} (median {sorted(ts)[2]:.3f})  max|diff vs lockstep| {float((out - ref).abs().max()):.3e}")
if flag & 16:
    import ctypes as C
    buf = (C.c_longlong * 8)()
    _lib.lib().dmnerf_debug_counters(buf)
    pairs = (n * s / 256) / 148
    for nm, v in zip(["issuer total", "issuer wait w_full", "issuer wait a_ready", "epi total", "epi wait acc_ready",
                      "producer wait w_empty"], buf):
        print(f"  {nm:24s} {v:12d}  per pair {v / pairs:10.0f}")
elif flag & 32:
    import ctypes as C
    buf = (C.c_longlong * 8)()
    _lib.lib().dmnerf_debug_counters(buf)
    duos = (n * s / 256) / 74
    for nm, v in zip(["issuer total", "issuer wait w_full", "epi signalling", "issuer wait a_ready", "epi total",
                      "epi wait acc_ready", "producer wait w_empty", "epi ldtm+wait"], buf):
        print(f"  {nm:24s} {v:12d}  per duo {v / duos:10.0f}")
elif flag & 8:
    import ctypes as C
    buf = (C.c_longlong * 8)()
    _lib.lib().dmnerf_debug_counters(buf)
    quads = (n * s / 512) / 74
    for nm, v in zip(["issuer total", "issuer wait w_full", "issuer wait peer_full", "issuer wait a_ready", "epi total",
                      "epi wait acc_ready", "producer wait w_empty"], buf):
        print(f"  {nm:24s} {v:12d}  per quad {v / quads:10.0f}")
