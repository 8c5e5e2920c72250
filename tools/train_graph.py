"""TrainStep eager vs CUDA-graph replay: step time and loss trajectory.  Usage: train_graph.py [N_rays] [K] [steps]"""
import os, sys, time
from types import SimpleNamespace
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
from dm_nerf_b200.train import TrainStep

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
k = int(sys.argv[2]) if len(sys.argv) > 2 else 32
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
ops.set_mode("bf16")
H, W = 480, 640
dev = torch.device("cuda")
pe, ve = get_embedder(10, 0)[0], get_embedder(4, 0)[0]
focal = 0.5 * W / np.tan(0.5 * 0.6911)
Kmat = torch.tensor([[focal, 0, W / 2], [0, -focal, H / 2], [0, 0, -1]], dtype=torch.float32)
pose = torch.eye(4)[:3, :4].clone(); pose[2, 3] = 7.0; pose = pose.to(dev)
img = torch.rand(H, W, 3, device=dev)
lab = ((torch.arange(H, device=dev)[:, None] // 60) * 8 + (torch.arange(W, device=dev)[None] // 80)) % k
targs = SimpleNamespace(N_train=n, near=4.0, far=15.0, N_samples=64, N_importance=128, perturb=1.0, ins_num=k,
                        tolerance=0.05, deta_w=0.05, lrate=5e-4, lrate_decay=500)
if os.environ.get("DMNERF_NCU"):
    # one eager, serial step between cudaProfilerStart/Stop (ncu --profile-from-start off)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
    ts = TrainStep(pe, ve, mc.to(dev), mf.to(dev), Kmat, targs, penalize=True, graph=False, overlap_assignment=False)
    for _ in range(3):
        ts.step(img, pose, lab)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    ts.step(img, pose, lab)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    sys.exit(0)
for graph, overlap in ((False, False), (True, False), (True, True)):
    torch.manual_seed(0)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
    mc, mf = mc.to(dev), mf.to(dev)
    ts = TrainStep(pe, ve, mc, mf, Kmat, targs, penalize=True, graph=graph, overlap_assignment=overlap)
    ts.reserved_sms = int(os.environ.get('DMNERF_RESERVE', ts.reserved_sms))
    w0 = mc.mlps[0].weight.detach().clone()
    losses = []
    for _ in range(5):
        losses.append(ts.step(img, pose, lab)["total"].clone())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        losses.append(ts.step(img, pose, lab)["total"].clone())
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    ls = [float(x) for x in losses]
    moved = float((mc.mlps[0].weight.detach() - w0).abs().max())
    print(f"graph={graph} overlap={overlap}: {dt*1e3:.2f} ms/step {n/dt:.0f} rays/s | loss first {ls[0]:.4f} mid {ls[len(ls)//2]:.4f} last {ls[-1]:.4f} "
          f"finite={all(np.isfinite(ls))} | max|dW0| {moved:.3e} | iteration {ts.iteration}")
if os.environ.get("DMNERF_PROFILE"):
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            ts.step(img, pose, lab)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=70))
