"""One training step of the reference's loop body (train_dmsr.py:24-72 / train_replica.py / train_scannet.py) on
the CUDA ray path, data-parallel over ray batches: pixel draw -> rays -> dm_nerf (coarse + fine) ->
MSE + Hungarian instance loss + emptiness penalizer -> backward -> ONE all-reduce of the flat gradient -> Adam
-> lr decay.  Host orchestration only: every kernel on the ray path is libdmnerf_b200.so's."""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Optional

import torch

from . import parallel
from .networks.evaluator import img2mse, ins_criterion_device, ins_criterion_multi
from .networks.helpers import get_select_full, z_val_sample
from .networks.penalizer import ins_penalizer
from .networks.render import dm_nerf


class TrainStep:
    """``step(image [H,W,3], pose [3,4], labels [H,W]) -> dict of detached loss scalars``.

    args: duck-typed like the reference's (N_train, near, far, N_samples, N_importance, perturb, ins_num, tolerance,
    deta_w, lrate, lrate_decay); ``penalize`` adds the emptiness loss (train_dmsr.py:53-60)."""

    def __init__(self, position_embedder, view_embedder, model_coarse, model_fine, K, args, penalize: bool = True,
                 optimizer: Optional[torch.optim.Optimizer] = None, device_assignment: bool = True, max_label: int = 256,
                 graph: bool = False):
        self.pe, self.ve, self.mc, self.mf, self.K = position_embedder, view_embedder, model_coarse, model_fine, K
        self.args = SimpleNamespace(**vars(args))
        self.args.is_train, self.args.N_ins = True, None
        self.penalize = penalize
        # device_assignment: Hungarian matching on the GPU (dmnerf_hungarian) -> no host round trip in the step;
        # False = scipy on the host as in the reference (evaluator.py:45-46), one synchronisation per step
        self.device_assignment, self.max_label = device_assignment, max_label
        self.params = list(model_coarse.parameters()) + list(model_fine.parameters())
        # graph=True: the whole step -- pixel draw, rays, both passes, losses, device Hungarian,
        # backward, gradient all-reduce, Adam -- is captured ONCE into a CUDA graph and replayed; image / pose / labels
        # are copied into static buffers and the learning rate lives in a device scalar.  Needs device_assignment.
        self.graph_mode = bool(graph)
        if self.graph_mode and not (device_assignment and optimizer is None):
            raise ValueError("TrainStep(graph=True) needs device_assignment=True and its own optimizer")
        # train_dmsr.py:127-128 builds torch.optim.Adam(params, lr, betas=(0.9, 0.999)); fused=True is the same update
        # in one launch for all 60 tensors instead of one foreach chain per step (0.6 -> 0.1 ms of host time)
        if self.graph_mode:
            self.lr_t = torch.tensor(float(args.lrate), device=self.params[0].device)
            self.opt = torch.optim.Adam(self.params, lr=self.lr_t, betas=(0.9, 0.999), fused=True, capturable=True)
        else:
            self.opt = optimizer or torch.optim.Adam(self.params, lr=args.lrate, betas=(0.9, 0.999),
                                                     fused=self.params[0].is_cuda)
        self._graph = self._static = self._static_out = None
        self._K_dev = None
        self.z_coarse = z_val_sample(args.N_train, args.near, args.far, args.N_samples)     # train_dmsr.py:22
        self.flat = None
        self.iteration = 0

    def step(self, image, pose, labels) -> Dict[str, torch.Tensor]:
        if self.graph_mode:
            return self._step_graph(image, pose, labels)
        return self._step_eager(image, pose, labels, self.K)

    def _step_graph(self, image, pose, labels) -> Dict[str, torch.Tensor]:
        a = self.args
        dev = self.params[0].device
        if self._static is None:
            self._K_dev = torch.as_tensor(self.K, dtype=torch.float32).reshape(3, 3).to(dev).contiguous()
            self._static = (torch.empty_like(image, device=dev), torch.empty((3, 4), dtype=torch.float32, device=dev),
                            torch.empty_like(labels, device=dev))
        s_img, s_pose, s_lab = self._static
        s_img.copy_(image, non_blocking=True)
        s_pose.copy_(torch.as_tensor(pose, dtype=torch.float32)[:3, :4], non_blocking=True)
        s_lab.copy_(labels, non_blocking=True)
        self.lr_t.fill_(a.lrate * (0.1 ** (self.iteration / (a.lrate_decay * 1000))))         # train_dmsr.py:68-72
        if self._graph is None:
            # warm-up on a side stream (allocator, lazy state of Adam, NCCL), then capture one step
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    self._step_eager(s_img, s_pose, s_lab, self._K_dev, decay=False)
            torch.cuda.current_stream().wait_stream(side)
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._static_out = self._step_eager(s_img, s_pose, s_lab, self._K_dev, decay=False)
            self.iteration += 3
        self._graph.replay()
        self.iteration += 1
        # the packed-weight caches were filled inside the graph: their keys do not see the replayed Adam updates
        self.mc._packed.clear(); self.mf._packed.clear()
        return self._static_out

    def _step_eager(self, image, pose, labels, K, decay: bool = True) -> Dict[str, torch.Tensor]:
        a = self.args
        target_c, target_i, rays = get_select_full(image, pose, K, labels, a.N_train)
        info = dm_nerf(rays, self.pe, self.ve, self.mc, self.mf, self.z_coarse, a)
        rgb_loss = img2mse(info["rgb_coarse"], target_c) + img2mse(info["rgb_fine"], target_c)
        emptiness = None
        if self.penalize:
            # enqueued BEFORE the instance loss so the device has work while the host runs the Hungarian assignment
            emptiness = ins_penalizer(info["raw_coarse"], info["z_vals_coarse"], info["depth_coarse"], rays[1], a) + \
                ins_penalizer(info["raw_fine"], info["z_vals_fine"], info["depth_fine"], rays[1], a)
        if self.device_assignment:
            ins_c, ins_f = ins_criterion_device([info["ins_coarse"], info["ins_fine"]], target_i, a.ins_num, self.max_label)
        else:
            ins_c, ins_f = ins_criterion_multi([info["ins_coarse"], info["ins_fine"]], target_i, a.ins_num)
        ins_loss = ins_c[0] + ins_f[0]
        total = rgb_loss + ins_loss
        if emptiness is not None:
            total = total + emptiness
        self.opt.zero_grad(set_to_none=True)          # backward then installs the fresh gradient tensors: no fills, no adds
        total.backward()
        if self.flat is None:
            self.flat = torch.empty(sum(p.numel() for p in self.params), dtype=torch.float32, device=self.params[0].device)
        parallel.allreduce_grads(self.params, self.flat)
        self.opt.step()
        if decay:
            lr = a.lrate * (0.1 ** (self.iteration / (a.lrate_decay * 1000)))              # train_dmsr.py:68-72
            for group in self.opt.param_groups:
                group["lr"] = lr
            self.iteration += 1
        out = {"total": total.detach(), "rgb": rgb_loss.detach(), "ins": ins_loss.detach()}
        if emptiness is not None:
            out["emptiness"] = emptiness.detach()
        return out
