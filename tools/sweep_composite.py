"""Launch-shape sweep of the bulk compositing kernels (ring depth x blocks per SM) through the debug-flag overrides."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_nerf_b200 import ops, _lib
from dm_nerf_b200._lib import COMP_DROP_EMPTY
n = 32768
dev = torch.device("cuda")
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timed(fn, iters=7):
    for _ in range(2): fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]
g = torch.Generator(device="cuda").manual_seed(0)
rd = torch.randn(n, 3, device=dev, generator=g)
for K in (32, 93):
    C = K + 5
    for S in (64, 192):
        raw = torch.randn(n, S, C, device=dev, generator=g)
        z = torch.sort(torch.rand(n, S, device=dev, generator=g) * 11 + 4, -1)[0]
        fb = n * (S * C * 4 + S * 4 + 12 + (3 + 1 + K) * 4 + S * 4)
        bb = n * (2 * S * C * 4 + S * 4 + 12 + (3 + 1 + K) * 4)
        rgb, w, depth, ins = ops.composite_fwd(raw, z, rd, COMP_DROP_EMPTY)
        d_rgb, d_depth, d_ins = torch.randn_like(rgb), torch.randn_like(depth), torch.randn_like(ins)
        for st in (0, 2, 3, 4):
            for ps in (0, 1, 2, 3, 4, 6):
                if st == 0 and ps != 0: continue
                _lib.lib().dmnerf_debug_flags(((st - 1) << 8 if st else 0) | (ps << 10))
                try:
                    f = timed(lambda: ops.composite_fwd(raw, z, rd, COMP_DROP_EMPTY))
                    b = timed(lambda: ops.composite_bwd(raw, z, rd, COMP_DROP_EMPTY, True, d_rgb, d_depth, None, d_ins))
                    print(f"K={K} S={S} stages={st} per_sm={ps}: fwd {f*1e3:7.1f} us {fb/f/1e6/peak:.3f} | bwd {b*1e3:7.1f} us {bb/b/1e6/peak:.3f}", flush=True)
                except Exception as e:
                    print(f"K={K} S={S} stages={st} per_sm={ps}: {str(e)[:80]}")
                    torch.cuda.synchronize()
        _lib.lib().dmnerf_debug_flags(0)
