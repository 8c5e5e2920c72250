import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dm_nerf_b200 import _lib
lib = _lib.lib()
out = torch.zeros(3, dtype=torch.int64, device="cuda")
for batched in (0, 1):
    for w in (4, 8, 16):
        lib.dmnerf_selftest_tmem_read_bw(w, 200, batched, C.c_void_p(out.data_ptr()), None)
        torch.cuda.synchronize()
        cyc, b = int(out[0]), int(out[1])
        print(f"warps={w:2d} batched={batched}: {cyc} cycles, {b / cyc:.1f} B/clk/SM, {cyc / 200:.0f} clk per 256 KB")
for n in (256, 128, 64, 48, 32):
    lib.dmnerf_selftest_umma_rate(n, 4096, C.c_void_p(out.data_ptr()), None)
    torch.cuda.synchronize()
    print(f"umma M=128 N={n:3d} K=16: {int(out[0]) / 4096:.1f} clk per MMA (floor 128*N/256 = {n // 2})")
