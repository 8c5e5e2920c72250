"""Times the emptiness-penalizer kernels alone: argv = n_rays K [mode] [depth_frac]."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dm_nerf_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
K = int(sys.argv[2]) if len(sys.argv) > 2 else 93
mode = sys.argv[3] if len(sys.argv) > 3 else "bf16"
frac = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
ops.set_mode(mode)
dev = "cuda"
for S in (64, 192):
    C = K + 5
    raw = torch.randn(n, S, C, device=dev)
    z = torch.sort(torch.rand(n, S, device=dev) * 6.5, -1)[0]
    rd = torch.randn(n, 3, device=dev)
    depth = torch.full((n,), 6.5 * frac, device=dev)
    flags = 0
    def run(f):
        for _ in range(3): f()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): f()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / 10 * 1e3
    loss, partials = ops.penalizer_fwd(raw, z, depth, rd, 0.05, 0.05, flags)
    g = torch.ones(1, device=dev)
    d_raw = torch.zeros_like(raw)
    t_f = run(lambda: ops.penalizer_fwd(raw, z, depth, rd, 0.05, 0.05, flags))
    t_b = run(lambda: ops.penalizer_bwd(raw, z, depth, rd, 0.05, 0.05, flags, partials, g, d_raw))
    t_a = run(lambda: ops.penalizer_bwd(raw, z, depth, rd, 0.05, 0.05, flags | 2, partials, g, d_raw))
    print(f"n={n} S={S} K={K} mode={mode} depth={frac}: fwd {t_f:.1f} us  bwd {t_b:.1f} us  bwd(accumulate) {t_a:.1f} us  "
          f"[{n*S*C*4/1e6:.0f} MB of logits]")
