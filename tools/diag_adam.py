"""Loss trajectory of TrainStep (eager) under three optimizers on identical seeds: FlatAdam, torch Adam, torch fused Adam."""
import os, sys
from types import SimpleNamespace
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dmnerf_oracle as orc
from dm_nerf_b200 import ops
from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
from dm_nerf_b200.train import TrainStep

n, k = 1024, 32
ops.set_mode(sys.argv[1] if len(sys.argv) > 1 else "bf16")
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
for name in ("flat", "torch", "torch_fused"):
    torch.manual_seed(0)
    mc, mf = DM_NeRF(8, 256, 63, 27, [4], k), DM_NeRF(8, 256, 63, 27, [4], k)
    mc.load_state_dict(orc.init_params(k, 1)); mf.load_state_dict(orc.init_params(k, 2))
    mc, mf = mc.to(dev), mf.to(dev)
    params = list(mc.parameters()) + list(mf.parameters())
    opt = None if name == "flat" else torch.optim.Adam(params, lr=5e-4, betas=(0.9, 0.999), fused=(name == "torch_fused"))
    ts = TrainStep(pe, ve, mc, mf, Kmat, targs, penalize=True, graph=False, optimizer=opt, overlap_assignment=False)
    w0 = mc.mlps[0].weight.detach().clone()
    ls = []
    for it in range(12):
        out = ts.step(img, pose, lab)
        ls.append(float(out["total"]))
        if it == 0:
            g0 = mc.mlps[0].weight.grad.detach().clone()
            print(name, "grad0 |max|", float(g0.abs().max()), "sum", float(g0.double().sum()))
    print(name, " ".join(f"{x:.4f}" for x in ls), "| max|dW0|", float((mc.mlps[0].weight.detach() - w0).abs().max()))
