"""exchange kernel variants on 16,384 rays x S, K=32: debug bit 8 = single-buffer kernel, default = double-buffered."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, _lib
n, k = 16384, 32
c = k + 5
gen = torch.Generator().manual_seed(0)
for s in (64, 320):
    ori = torch.randn(n, s, c, generator=gen).cuda(); tar = torch.randn(n, s, c, generator=gen).cuda()
    oa = torch.rand(n, k + 1, generator=gen).cuda(); ta = torch.rand(n, k + 1, generator=gen).cuda()
    outs = {}
    for flag in (0, 0):
        _lib.lib().dmnerf_debug_flags(flag)
        o = ori.clone()
        ops.exchange(o, tar, oa, ta, 8, None, want_ori_label=False, want_tar_label=False)
        torch.cuda.synchronize()
        work = [ori.clone() for _ in range(6)]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for w in work: ops.exchange(w, tar, oa, ta, 8, None, want_ori_label=False, want_tar_label=False)
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / len(work)
        outs[flag] = o
        gb = 2 * n * s * c * 4 / 1e9
        print(f"S={s} flag {flag}: {ms * 1000:.1f} us  reads {gb / ms * 1000:.0f} GB/s  rows changed {float((o != ori).any(-1).float().mean()):.3f}", flush=True)
    
_lib.lib().dmnerf_debug_flags(0)
