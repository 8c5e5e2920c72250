"""Per-ring-step cost paid by the MMA-issuing thread (dmnerf_selftest_issue_cost)."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dm_nerf_b200 import _lib
lib = _lib.lib()
out = torch.zeros(2, dtype=torch.int64, device="cuda")
names = {0: "2 MMAs", 1: "+commit", 2: "+wait", 4: "+fence", 8: "+clock64", 3: "+commit+wait", 7: "+commit+wait+fence", 15: "all", 16: "elect: 2 MMAs", 17: "elect +commit", 19: "elect +commit+wait", 23: "elect +commit+wait+fence", 48: "elect acc=1", 80: "elect 4 MMAs", 112: "elect 4 MMAs acc=1", 240: "elect 4 MMAs acc=1 2 accumulators", 243: "same +commit+wait"}
for mode, nm in names.items():
    for _ in range(2):
        lib.dmnerf_selftest_issue_cost(mode, 4096, C.c_void_p(out.data_ptr()), None)
        torch.cuda.synchronize()
    print(f"mode {mode:2d} {nm:22s} {out[0].item() / 4096:8.1f} clk per group")
