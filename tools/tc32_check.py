"""Error of the tensor-core modes against the fp32 CUDA-core mode on raw MLP outputs, and their launch times
(32768 rays x 192 samples, K=32).  Run on the GPU box: python tools/tc32_check.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dm_nerf_b200 import ops, synth
from dm_nerf_b200.networks.dm_nerf import DM_NeRF

dev = "cuda"
k = 32
m = DM_NeRF(8, 256, 63, 27, [4], k)
m.load_state_dict(synth.init_state_dict(k, 11, float(os.environ.get("GAIN", "1.0"))))
m = m.to(dev)
gen = torch.Generator().manual_seed(4)
for n, s in ((3, 77), (40, 192), (1000, 192)):
    o = ((torch.rand(n, 3, generator=gen) - 0.5) * 2).to(dev)
    d = torch.randn(n, 3, generator=gen).to(dev)
    z = torch.sort(torch.rand(n, s, generator=gen) * 11 + 4, -1)[0].to(dev)
    with torch.no_grad():
        ops.set_mode("fp32"); ref = m.query_rays(o, d, z)
        for mode in ("bf16", "fp16", "tc32"):
            ops.set_mode(mode); got = m.query_rays(o, d, z)
            e = (got - ref).abs()
            print(f"n={n} s={s} {mode}: max {float(e.max()):.3e} mean {float(e.mean()):.3e}  (|ref| max {float(ref.abs().max()):.2f})", flush=True)
n, s = 32768, 192
o = ((torch.rand(n, 3, generator=gen) - 0.5) * 2).to(dev)
d = torch.randn(n, 3, generator=gen).to(dev)
z = torch.sort(torch.rand(n, s, generator=gen) * 11 + 4, -1)[0].to(dev)
for mode in ("bf16", "fp16", "tc32", "bf16", "fp16", "tc32"):
    ops.set_mode(mode)
    with torch.no_grad():
        for _ in range(2): m.query_rays(o, d, z)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): m.query_rays(o, d, z)
        b.record(); torch.cuda.synchronize()
    print(f"{mode}: {a.elapsed_time(b) / 5:.3f} ms per 32768x192", flush=True)
# cycle accounting of block 0 (debug bit 4)
import ctypes as C
from dm_nerf_b200 import _lib
ops.set_mode("tc32")
_lib.lib().dmnerf_debug_flags(16)
with torch.no_grad():
    m.query_rays(o, d, z)
buf = (C.c_longlong * 12)()
_lib.lib().dmnerf_debug_counters_tc32(buf)
_lib.lib().dmnerf_debug_flags(0)
tiles = (n * s / 128) / 148
names = ["issuer total", "issuer wait w_full", "issuer wait a_ready", "-", "-", "-",
         "-", "-", "epi total", "epi wait acc_ready", "producer wait w_empty", "-"]
for nm, v in zip(names, buf):
    if nm != "-":
        print(f"  {nm:24s} {v:12d}  per tile {v / tiles:10.0f}")
