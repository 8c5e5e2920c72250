"""The optimizer of the reference's training loop -- ``torch.optim.Adam(params, lr, betas=(0.9, 0.999))``
(train_dmsr.py:127-128, stepped at :62-66) -- over ONE flat fp32 buffer per model.

The parameters of each DM_NeRF become views into a flat buffer laid out exactly like the flat gradient buffer the MLP
backward writes (ops.mlp_bwd: state_dict order, every tensor padded to 4 floats), so a step is one
``dmnerf_adam_step`` launch per model on (p, g, m, v) with no gather/scatter, the data-parallel all-reduce runs on the
gradient buffer in place, and its 1/world mean is folded into the update.  Step count and learning rate live in device
scalars: the step can sit in a CUDA graph."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib
from .ops import _ptr, _stream, check, _count


class FlatAdam:
    def __init__(self, models: Sequence[torch.nn.Module], lr: float, betas=(0.9, 0.999), eps: float = 1e-8,
                 lr_tensor: Optional[torch.Tensor] = None):
        self.models = list(models)
        self.lr, self.betas, self.eps, self.lr_tensor = float(lr), (float(betas[0]), float(betas[1])), float(eps), lr_tensor
        self.groups: List[Dict] = []
        for mod in self.models:
            params = mod._ordered_params()
            dev = params[0].device
            if not params[0].is_cuda:
                raise RuntimeError("FlatAdam: parameters must live on a CUDA device (dm_nerf_b200 has no CPU path)")
            sizes = [p.numel() for p in params]
            padded = [(n + 3) & ~3 for n in sizes]
            offs = [0]
            for n in padded:
                offs.append(offs[-1] + n)
            flat = torch.zeros(offs[-1], dtype=torch.float32, device=dev)
            for p, o, n in zip(params, offs, sizes):
                view = flat[o:o + n].view(p.shape)
                view.copy_(p.data)
                p.data = view                                   # same values, now a window of the flat buffer
            self.groups.append({"module": mod, "params": params, "offsets": offs[:-1], "sizes": sizes, "total": offs[-1],
                                "p": flat, "m": torch.zeros_like(flat), "v": torch.zeros_like(flat), "gather": None})
            mod._packed.clear()
        self.step_t = torch.zeros((), dtype=torch.int32, device=self.groups[0]["p"].device)
        self.params = [p for g in self.groups for p in g["params"]]

    # -- torch.optim surface used by the loop -------------------------------------------------------------------
    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if set_to_none or p.grad is None:
                p.grad = None
            else:
                p.grad.zero_()

    def grad_buffers(self) -> List[torch.Tensor]:
        """One flat gradient tensor per model, in the parameter buffer's layout.  The MLP backward hands autograd
        views of one zero-initialised flat buffer, which is then used in place; gradients that arrived any other way
        are gathered into a buffer of the same layout."""
        return [self.grad_buffer(i) for i in range(len(self.groups))]

    def grad_buffer(self, i: int) -> torch.Tensor:
        """The flat gradient tensor of model ``i`` alone (ready as soon as THAT model's backward has run)."""
        for g in (self.groups[i],):
            grads = [p.grad for p in g["params"]]
            if any(x is None for x in grads):
                raise RuntimeError("FlatAdam: a parameter has no gradient (call backward() first)")
            st, base = grads[0].untyped_storage(), grads[0].storage_offset()
            in_place = grads[0].dtype == torch.float32 and st.nbytes() >= 4 * (base + g["total"]) and all(
                x.untyped_storage().data_ptr() == st.data_ptr() and x.storage_offset() == base + o and x.is_contiguous()
                for x, o in zip(grads, g["offsets"]))
            if in_place:
                flat = torch.empty(0, dtype=torch.float32, device=grads[0].device).set_(st, base, (g["total"],))
            else:
                if g["gather"] is None:
                    g["gather"] = torch.zeros(g["total"], dtype=torch.float32, device=g["p"].device)
                flat = g["gather"]
                torch._foreach_copy_([flat[o:o + n] for o, n in zip(g["offsets"], g["sizes"])], [x.reshape(-1) for x in grads])
            return flat

    def step(self, grad_buffers: Optional[List[torch.Tensor]] = None, grad_scale: float = 1.0) -> None:
        bufs = grad_buffers if grad_buffers is not None else self.grad_buffers()
        self.step_t += 1
        lib = _lib.lib()
        for g, gb in zip(self.groups, bufs):
            check(lib.dmnerf_adam_step(_ptr(g["p"]), _ptr(gb), _ptr(g["m"]), _ptr(g["v"]), g["total"], _ptr(self.lr_tensor),
                                       C.c_float(self.lr), _ptr(self.step_t), C.c_float(self.betas[0]), C.c_float(self.betas[1]),
                                       C.c_float(self.eps), C.c_float(grad_scale), _stream()), "dmnerf_adam_step")
            g["module"]._packed.clear()            # the kernel writes through raw pointers: no version-counter bump
        _count(len(self.groups))

    # -- checkpoint compatibility (train_dmsr.py:80-88 saves optimizer.state_dict()) ---------------------------------
    def state_dict(self) -> Dict:
        """torch.optim.Adam's format, so a checkpoint written from here loads into the reference's optimizer."""
        state, idx = {}, 0
        for g in self.groups:
            for o, n, p in zip(g["offsets"], g["sizes"], g["params"]):
                state[idx] = {"step": self.step_t.detach().to(torch.float32).clone(),
                              "exp_avg": g["m"][o:o + n].view(p.shape).clone(),
                              "exp_avg_sq": g["v"][o:o + n].view(p.shape).clone()}
                idx += 1
        lr = float(self.lr_tensor) if self.lr_tensor is not None else self.lr
        return {"state": state, "param_groups": [{"lr": lr, "betas": self.betas, "eps": self.eps, "weight_decay": 0,
                                                  "amsgrad": False, "maximize": False, "params": list(range(idx))}]}

    def load_state_dict(self, sd: Dict) -> None:
        idx = 0
        for g in self.groups:
            for o, n, p in zip(g["offsets"], g["sizes"], g["params"]):
                st = sd["state"].get(idx)
                if st is not None:
                    g["m"][o:o + n].copy_(st["exp_avg"].reshape(-1))
                    g["v"][o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
                    self.step_t.fill_(int(float(st["step"])))
                idx += 1
        if sd.get("param_groups"):
            self.lr = float(sd["param_groups"][0]["lr"])
            if self.lr_tensor is not None:               # the value a captured step actually reads
                self.lr_tensor.fill_(self.lr)
