"""One training step of the reference's loop body (train_dmsr.py:24-72 / train_replica.py / train_scannet.py) on
the CUDA ray path, data-parallel over ray batches: pixel draw -> rays -> dm_nerf (coarse + fine) ->
MSE + Hungarian instance loss + emptiness penalizer -> backward -> ONE all-reduce of the flat gradient -> Adam
-> lr decay.  Host orchestration only: every kernel on the ray path is libdmnerf_b200.so's."""
from __future__ import annotations

from types import SimpleNamespace
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import parallel
from .optim import FlatAdam
from . import ops
from .networks.evaluator import (img2mse, ins_criterion_device, ins_criterion_multi, ins_criterion_prepped, ins_label_prep,
                                 ins_loss_and_grad_fused)
from .networks.helpers import get_select_crop, get_select_full, z_val_sample
from .networks.render import dm_nerf


class TrainStep:
    """``step(image [H,W,3], pose [3,4], labels [H,W]) -> dict of detached loss scalars``.

    args: duck-typed like the reference's (N_train, near, far, N_samples, N_importance, perturb, ins_num, tolerance,
    deta_w, lrate, lrate_decay); ``penalize`` adds the emptiness loss (train_dmsr.py:53-60)."""

    def __init__(self, position_embedder, view_embedder, model_coarse, model_fine, K, args, penalize: bool = True,
                 optimizer: Optional[torch.optim.Optimizer] = None, device_assignment: bool = True, max_label: int = 256,
                 graph: bool = False, overlap_assignment: Optional[bool] = None):
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
        # train_dmsr.py:127-128 builds torch.optim.Adam(params, lr, betas=(0.9, 0.999)).  Default here: the same update
        # as one dmnerf_adam_step launch per model over flat buffers (optim.py; torch's fused multi-tensor Adam took
        # 0.155 ms per step for the 60 tensors); pass ``optimizer`` to keep a torch optimizer (eager mode only).
        self.lr_t = torch.tensor(float(args.lrate), device=self.params[0].device) if self.graph_mode else None
        self.opt = optimizer or FlatAdam([model_coarse, model_fine], lr=args.lrate, betas=(0.9, 0.999), lr_tensor=self.lr_t)
        # overlap_assignment (default: on with graph=True): the two Hungarian assignments are serial one-warp kernels
        # (~1 ms at K=93), so they run on side streams -- the coarse one beside the fine pass, the fine one beside the
        # coarse model's backward (the two models' autograd graphs are disjoint: render.py:68 detaches the samples) --
        # while the persistent MLP kernels leave `reserved_sms` SMs free for them (dmnerf_reserve_sms).
        self.overlap = bool(self.graph_mode if overlap_assignment is None else overlap_assignment) and device_assignment
        self.reserved_sms = 1
        # the instance loss and its gradient as five launches (ops.ins_assign_loss_grad) instead of ~130 tensor ops +
        # autograd.grad on the side streams; False keeps the tensor-op formulation (same values, tested)
        self.fused_ins_loss = True
        self._side = None
        self._early_reduce = {}                # model index -> (flat gradient, pending all-reduce) started before _finish
        self._draw_step = self._draw_seed = None
        self._graph = self._static = self._static_out = None
        self._K_dev = None
        self.z_coarse = z_val_sample(args.N_train, args.near, args.far, args.N_samples)     # train_dmsr.py:22
        self.flat = None
        self.iteration = 0

    def step(self, image, pose, labels, ins_index=None, crop_mask=None) -> Dict[str, torch.Tensor]:
        """``ins_index`` / ``crop_mask``: the ScanNet loop (train_scannet.py:24-66) -- pixels drawn by ``get_select_crop``
        (30 % labelled, placed last), instance losses on the last ``N_ins`` rays only (render.py:88-90).  That draw is
        the reference's host-side numpy code, so this variant runs eagerly and serially."""
        if ins_index is not None:
            if self.graph_mode:
                raise ValueError("TrainStep(graph=True) cannot run the ScanNet step: get_select_crop draws on the host")
            return self._step_serial(image, pose, labels, self.K, crop=(ins_index, crop_mask))
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
        # step i runs with lr(i-1), as the eager path and the reference do: the decay is applied AFTER the optimizer
        # step (train_dmsr.py:66-72), so it first acts on the next iteration
        self.lr_t.fill_(a.lrate * (0.1 ** (max(self.iteration - 1, 0) / (a.lrate_decay * 1000))))
        if self._graph is None:
            # warm-up on a side stream (allocator, NCCL communicators, lazy module state), then capture one step.  The
            # warm-up steps are real optimizer steps on this first batch, so they run on throw-away state: parameters,
            # Adam moments, step counter and the pixel-draw counter are snapshotted and put back after the capture --
            # the first replay is then step 0 of the schedule, exactly like the eager path.
            if self._draw_step is None:
                self._draw(s_img)                             # creates the device counter; the draw itself is discarded
                self._draw_step.zero_()
            snap = [(g["p"].clone(), g["m"].clone(), g["v"].clone()) for g in self.opt.groups]
            snap_step, snap_draw = self.opt.step_t.clone(), self._draw_step.clone()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    self._step_eager(s_img, s_pose, s_lab, self._K_dev, decay=False)
            torch.cuda.current_stream().wait_stream(side)
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._static_out = self._step_eager(s_img, s_pose, s_lab, self._K_dev, decay=False)
            for g, (p0, m0, v0) in zip(self.opt.groups, snap):
                g["p"].copy_(p0); g["m"].copy_(m0); g["v"].copy_(v0)
            self.opt.step_t.copy_(snap_step)
            self._draw_step.copy_(snap_draw)
            self.mc._packed.clear(); self.mf._packed.clear()
        self._graph.replay()
        self.iteration += 1
        # the packed-weight caches were filled inside the graph: their keys do not see the replayed Adam updates
        self.mc._packed.clear(); self.mf._packed.clear()
        return self._static_out

    def _step_eager(self, image, pose, labels, K, decay: bool = True) -> Dict[str, torch.Tensor]:
        if self.overlap:
            prev = ops.reserve_sms(self.reserved_sms)
            try:
                return self._step_overlapped(image, pose, labels, K, decay)
            finally:
                ops.reserve_sms(prev)
        return self._step_serial(image, pose, labels, K, decay)

    def _draw(self, image) -> torch.Tensor:
        """N_train distinct pixels in one launch (keyed permutation of the H*W indices, key = seed x DEVICE step counter, so
        a graph replay draws anew) instead of torch.randperm's radix sort of all H*W keys every step."""
        if self._draw_step is None:
            self._draw_step = torch.zeros((), dtype=torch.int32, device=image.device)
            self._draw_seed = int(torch.initial_seed()) * 1000003 + (dist.get_rank() if dist.is_initialized() else 0)
        self._draw_step += 1
        return ops.draw_pixels(int(self.args.N_train), int(image.shape[0]) * int(image.shape[1]), self._draw_step, self._draw_seed)

    def _finish(self, decay: bool) -> None:
        if isinstance(self.opt, FlatAdam):
            # ONE all-reduce per model on the backward's own flat gradient buffer (no gather / scatter copies); the
            # 1/world of the mean is applied inside the update
            world = dist.get_world_size() if dist.is_initialized() else 1
            early, self._early_reduce = self._early_reduce, {}
            bufs = [early[i][0] if i in early else self.opt.grad_buffer(i) for i in range(len(self.opt.groups))]
            if world > 1:
                for i, b in enumerate(bufs):
                    if i in early:
                        early[i][1].wait()                    # started right after that model's backward (see below)
                    else:
                        dist.all_reduce(b, op=dist.ReduceOp.SUM)
            self.opt.step(bufs, grad_scale=1.0 / world)
        else:
            if self.flat is None:
                self.flat = torch.empty(sum(p.numel() for p in self.params), dtype=torch.float32, device=self.params[0].device)
            parallel.allreduce_grads(self.params, self.flat)
            self.opt.step()
            # torch's fused multi-tensor Adam updates the parameters without moving their version counters, which are
            # the packed-weight caches' keys: without this the next forward would run on stale operand images
            self.mc._packed.clear(); self.mf._packed.clear()
        if decay:
            a = self.args
            lr = a.lrate * (0.1 ** (self.iteration / (a.lrate_decay * 1000)))              # train_dmsr.py:68-72
            if isinstance(self.opt, FlatAdam):
                self.opt.lr = lr
            else:
                for group in self.opt.param_groups:
                    group["lr"] = lr
            self.iteration += 1

    def _step_overlapped(self, image, pose, labels, K, decay: bool) -> Dict[str, torch.Tensor]:
        a = self.args
        main = torch.cuda.current_stream()
        if self._side is None:
            self._side = (torch.cuda.Stream(), torch.cuda.Stream())
        side_c, side_f = self._side
        # the operand images of the updated weights are packed on a side stream while the main stream draws the pixels
        # and prepares the label one-hot (about 0.1 ms of small kernels on each side)
        side_c.wait_stream(main)
        with torch.cuda.stream(side_c):
            self.mc.packed(ops.get_mode()); self.mf.packed(ops.get_mode())
        target_c, target_i, rays = get_select_full(image, pose, K, labels, a.N_train, idx=self._draw(image))
        main.wait_stream(side_c)
        # the label side of the instance loss (one-hot slots, their sort, counts: ~20 small kernels) is not needed before
        # the coarse pass is composited: it runs on the coarse side stream, beside the coarse forward
        side_c.wait_stream(main)
        with torch.cuda.stream(side_c):
            prep = ins_label_prep(target_i, a.ins_num, self.max_label)
            prep_done = torch.cuda.Event()
            prep_done.record(side_c)
        got = {}

        def ins_loss_and_grad(ins_map):
            # the instance loss AND its gradient w.r.t. the composited map, both on the side stream: the ~100 small
            # kernels of that backward then run beside the ray kernels instead of in front of the compositing backward
            if self.fused_ins_loss:
                loss, _, grad = ins_loss_and_grad_fused(ins_map, prep)
                return loss, grad
            with torch.enable_grad():                  # (the fine hook runs inside an autograd node's forward)
                leaf = ins_map.detach().requires_grad_(True)
                loss = ins_criterion_prepped([leaf], prep)[0][0]
                (grad,) = torch.autograd.grad(loss, leaf)
            return loss.detach(), grad

        def on_coarse(ins_coarse):
            side_c.wait_stream(main)
            with torch.cuda.stream(side_c):
                got["c"] = ins_loss_and_grad(ins_coarse)

        def on_fine(ins_fine):
            # ... and the fine assignment the moment the fine pass is composited, ahead of the fine penalizer launch: its
            # chain (label-wise sums, the one-warp Hungarian, ~100 small kernels of loss gradient) is longer than the
            # coarse model's backward it runs beside
            side_f.wait_stream(main)
            side_f.wait_event(prep_done)
            with torch.cuda.stream(side_f):
                got["f"] = ins_loss_and_grad(ins_fine)

        pen = (a.tolerance, a.deta_w) if self.penalize else None     # emptiness loss inside the compositing nodes
        info = dm_nerf(rays, self.pe, self.ve, self.mc, self.mf, self.z_coarse, a, on_coarse=on_coarse, penalizer=pen,
                       on_fine=on_fine)
        ins_f, g_ins_f = got["f"]
        # coarse model: losses and backward on the main stream while the fine assignment runs on its side stream
        rgb_c = img2mse(info["rgb_coarse"], target_c)
        pen_c = info.get("emptiness_coarse")
        main.wait_stream(side_c)
        ins_c, g_ins_c = got["c"]
        part_c = rgb_c if pen_c is None else rgb_c + pen_c
        self.opt.zero_grad(set_to_none=True)
        torch.autograd.backward([part_c, info["ins_coarse"]], [None, g_ins_c])
        # the coarse model's gradient is complete here (render.py:68 detaches the fine samples): its all-reduce starts
        # now, on NCCL's stream, and hides under the three-times-larger fine backward
        if isinstance(self.opt, FlatAdam) and dist.is_initialized() and dist.get_world_size() > 1:
            buf = self.opt.grad_buffer(0)
            self._early_reduce[0] = (buf, dist.all_reduce(buf, op=dist.ReduceOp.SUM, async_op=True))
        rgb_f = img2mse(info["rgb_fine"], target_c)
        pen_f = info.get("emptiness_fine")
        main.wait_stream(side_f)
        part_f = rgb_f if pen_f is None else rgb_f + pen_f
        torch.autograd.backward([part_f, info["ins_fine"]], [None, g_ins_f])
        self._finish(decay)
        rgb_loss, ins_loss = (rgb_c + rgb_f).detach(), ins_c + ins_f
        out = {"total": (part_c + part_f).detach().reshape(()) + ins_loss, "rgb": rgb_loss, "ins": ins_loss}
        if self.penalize:
            out["emptiness"] = (pen_c + pen_f).detach()
        return out

    def _step_serial(self, image, pose, labels, K, decay: bool = True, crop=None) -> Dict[str, torch.Tensor]:
        a = self.args
        if crop is not None:
            a = SimpleNamespace(**vars(a))
            target_c, target_i, rays, a.N_ins = get_select_crop(image, pose, K, labels, crop[0], crop[1], a.N_train)
        else:
            target_c, target_i, rays = get_select_full(image, pose, K, labels, a.N_train, idx=self._draw(image))
        pen = (a.tolerance, a.deta_w) if self.penalize else None     # emptiness loss inside the compositing nodes
        info = dm_nerf(rays, self.pe, self.ve, self.mc, self.mf, self.z_coarse, a, penalizer=pen)
        rgb_loss = img2mse(info["rgb_coarse"], target_c) + img2mse(info["rgb_fine"], target_c)
        emptiness = info["emptiness_coarse"] + info["emptiness_fine"] if self.penalize else None
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
        self._finish(decay)
        out = {"total": total.detach(), "rgb": rgb_loss.detach(), "ins": ins_loss.detach()}
        if emptiness is not None:
            out["emptiness"] = emptiness.detach()
        return out
