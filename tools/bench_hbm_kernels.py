"""CUDA-event timing of the HBM-bound ray kernels (compositing fwd/bwd, sampler, penalizer fwd/bwd) against the
measured HBM peak, with SURVEY 8(d)'s algorithmic bytes.  Usage: bench_hbm_kernels.py [n_rays] [K]"""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dm_nerf_b200 import ops
from dm_nerf_b200._lib import COMP_DROP_EMPTY

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
K = int(sys.argv[2]) if len(sys.argv) > 2 else 32
C = K + 5
dev = torch.device("cuda")
peak = 6546.6
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


g = torch.Generator(device="cuda").manual_seed(0)
rd = torch.randn(n, 3, device=dev, generator=g)
for S in (64, 192):
    raw = torch.randn(n, S, C, device=dev, generator=g)
    z = torch.sort(torch.rand(n, S, device=dev, generator=g) * 11 + 4, -1)[0]
    fwd_bytes = n * (S * C * 4 + S * 4 + 12 + (3 + 1 + K) * 4 + S * 4)
    ms = timed(lambda: ops.composite_fwd(raw, z, rd, COMP_DROP_EMPTY))
    print(f"composite_fwd S={S} K={K}: {ms*1e3:8.1f} us  {fwd_bytes/ms/1e6:7.0f} GB/s  {fwd_bytes/ms/1e6/peak:.3f} of measured HBM peak")
    rgb, w, depth, ins = ops.composite_fwd(raw, z, rd, COMP_DROP_EMPTY)
    d_rgb, d_depth, d_ins = torch.randn_like(rgb), torch.randn_like(depth), torch.randn_like(ins)
    bwd_bytes = n * (2 * S * C * 4 + S * 4 + 12 + (3 + 1 + K) * 4)
    ms = timed(lambda: ops.composite_bwd(raw, z, rd, COMP_DROP_EMPTY, True, d_rgb, d_depth, None, d_ins))
    print(f"composite_bwd S={S} K={K}: {ms*1e3:8.1f} us  {bwd_bytes/ms/1e6:7.0f} GB/s  {bwd_bytes/ms/1e6/peak:.3f} (recompute + reverse pass)")
    _, w5, _, ins5, t5 = ops.composite_fwd(raw, z, rd, COMP_DROP_EMPTY, want_trans=True)
    ms = timed(lambda: ops.composite_bwd(raw, z, rd, COMP_DROP_EMPTY, True, d_rgb, d_depth, None, d_ins, fwd=(w5, t5, ins5)))
    rb = bwd_bytes + n * 2 * S * 4
    print(f"composite_bwd S={S} K={K}: {ms*1e3:8.1f} us  {rb/ms/1e6:7.0f} GB/s  {rb/ms/1e6/peak:.3f} (one reverse pass on the forward's w, T)")
    pen_bytes = n * (S * (K + 1) * 4 + S * 4 + 16)
    rawg = raw.clone().requires_grad_(True)
    def pen_f():
        return ops.penalizer(rawg, z, depth, rd, 0.05, 0.05)
    ms = timed(pen_f)
    print(f"penalizer_fwd S={S} K={K}: {ms*1e3:8.1f} us  {pen_bytes/ms/1e6:7.0f} GB/s  {pen_bytes/ms/1e6/peak:.3f}")
    loss = pen_f()
    def pen_b():
        rawg.grad = None
        loss.backward(retain_graph=True)
    ms = timed(pen_b)
    print(f"penalizer_bwd S={S} K={K}: {ms*1e3:8.1f} us  {(pen_bytes + n*S*C*4)/ms/1e6:7.0f} GB/s  {(pen_bytes + n*S*C*4)/ms/1e6/peak:.3f} (d_raw written with all C channels)")
zc = torch.sort(torch.rand(n, 64, device=dev, generator=g) * 11 + 4, -1)[0]
wc = torch.rand(n, 64, device=dev, generator=g)
ms = timed(lambda: ops.fine_z(zc, wc, 128, None))
print(f"fine_z det: {ms*1e3:8.1f} us  {n*1280/ms/1e6:7.0f} GB/s  {n*1280/ms/1e6/peak:.3f}")
u = torch.rand(n, 128, device=dev, generator=g)
ms = timed(lambda: ops.fine_z(zc, wc, 128, u))
print(f"fine_z rand: {ms*1e3:8.1f} us  {n*1792/ms/1e6:7.0f} GB/s  {n*1792/ms/1e6/peak:.3f}")
# exchanger (a10): one move label on [n, S, C] fields; SURVEY 8d bytes per (ray, target)
for S in (64, 192):
    ori = torch.randn(n, S, C, device=dev, generator=g) * 3
    tar = torch.randn(n, S, C, device=dev, generator=g) * 3
    oacc = torch.sigmoid(torch.randn(n, K + 1, device=dev, generator=g) * 3)
    tacc = torch.sigmoid(torch.randn(n, K + 1, device=dev, generator=g) * 3)
    ex_bytes = n * (2 * S * C * 4 + 2 * (K + 1) * 4 + S * C * 4 + 2 * S * 8)
    ms = timed(lambda: ops.exchange(ori, tar, oacc, tacc, 3, None))
    print(f"exchanger S={S} K={K}: {ms*1e3:8.1f} us  {ex_bytes/ms/1e6:7.0f} GB/s  {ex_bytes/ms/1e6/peak:.3f} (reads both fields, "
          f"rewrites ori_raw, writes two int64 label maps)")
