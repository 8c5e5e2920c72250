import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dm_nerf_b200 import ops
for k in (32, 93, 123):
    s = torch.rand(2, k, k, device="cuda") * 3
    nv = torch.tensor(k, dtype=torch.int32, device="cuda")
    for _ in range(3): ops.hungarian(s, nv)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): ops.hungarian(s, nv)
    b.record(); torch.cuda.synchronize()
    print(f"hungarian K={k}: {a.elapsed_time(b)/10*1e3:.1f} us")
