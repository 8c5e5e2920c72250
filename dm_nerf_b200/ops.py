"""Tensor-level wrappers over the C-ABI (one function per entry point) plus the autograd
Functions that give the ray path its backward.  torch is used for device memory, streams and
autograd bookkeeping only; every computation happens in libdmnerf_b200.so.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import MODE_BF16, MODE_F16, MODE_FP32, MODE_TC32, COMP_DROP_EMPTY, COMP_RAW_INS, NUM_PARAMS, check

_MODE_NAMES = {"fp32": MODE_FP32, "bf16": MODE_BF16, "fp16": MODE_F16, "f16": MODE_F16, "tc32": MODE_TC32}
_default_mode = _MODE_NAMES[os.environ.get("DMNERF_MODE", "fp32").lower()]

# launch counter: how many of OUR kernels-launching C-ABI calls were made (bench reports it)
launches = 0


def set_mode(mode) -> None:
    """Arithmetic of the MLP queries: 'fp32' (CUDA cores, parity mode), 'bf16' (tcgen05 chain, forward and -- with
    gradients enabled -- backward: the training mode), 'fp16' (same chain, fp16 operands, inference only: the rendering
    mode), 'tc32' (tcgen05, fp16 hi+lo split x3: fp32-accurate, inference only)."""
    global _default_mode
    _default_mode = _MODE_NAMES[mode] if isinstance(mode, str) else int(mode)


def get_mode() -> int:
    return _default_mode


def variant_selected() -> bool:
    """True while a parked forward variant is selected by the debug flags (their operand streams are packed on demand)."""
    return bool(_lib.lib().dmnerf_debug_flags(-1) & (4 | 8 | 32))


def reserve_sms(n: int) -> int:
    """SMs the persistent MLP kernels leave free for kernels on other streams (``dmnerf_reserve_sms``); returns the
    previous value.  Used by the overlapped training step; 0 everywhere else."""
    return int(_lib.lib().dmnerf_reserve_sms(int(n)))


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: tensor must live on a CUDA device (dm_nerf_b200 has no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _ptr(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _count(n: int = 1) -> None:
    global launches
    launches += n


class KernelTimer:
    """Per-kernel device timing with CUDA events on the launching (current) stream.  Usage:
    ``with KernelTimer() as kt: ...`` then ``kt.totals()`` -> {name: (launches, total_ms)}."""

    active = None

    def __init__(self):
        self.records = []

    def __enter__(self):
        KernelTimer.active = self
        return self

    def __exit__(self, *exc):
        KernelTimer.active = None

    def totals(self):
        torch.cuda.synchronize()
        out = {}
        for name, a, b in self.records:
            n, ms = out.get(name, (0, 0.0))
            out[name] = (n + 1, ms + a.elapsed_time(b))
        return out


def _timed_call(name: str, fn, *args) -> None:
    kt = KernelTimer.active
    if kt is None:
        check(fn(*args), name)
        return
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    check(fn(*args), name)
    b.record()
    kt.records.append((name, a, b))


# ---------------------------------------------------------------------------------------------------
# weights
# ---------------------------------------------------------------------------------------------------

PARAM_ORDER: Tuple[str, ...] = tuple(
    [f"mlps.{i}.{w}" for i in range(8) for w in ("weight", "bias")]
    + [f"{m}.{w}" for m in ("rgb_feature_linear", "ins_feature_linear", "rgb_feature_linears.0",
                           "ins_feature_linears.0", "density_linear", "ins_linear", "rgb_linear")
       for w in ("weight", "bias")])
assert len(PARAM_ORDER) == NUM_PARAMS


def _ptr_array(tensors: Sequence[torch.Tensor]):
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def pack_weights(params: Sequence[torch.Tensor], ins_num: int, mode: int) -> torch.Tensor:
    """params: the 30 tensors in state_dict order (fp32, CUDA).  Returns the packed image (uint8)."""
    params = [_f32(p.detach(), "param") for p in params]
    nbytes = _lib.lib().dmnerf_packed_bytes(ins_num, mode)
    if nbytes <= 0:
        raise RuntimeError(f"pack_weights: unsupported ins_num={ins_num} / mode={mode}: {_lib.last_error()}")
    out = torch.empty(nbytes, dtype=torch.uint8, device=params[0].device)
    check(_lib.lib().dmnerf_weights_pack(_ptr_array(params), ins_num, mode, _ptr(out), _stream()), "dmnerf_weights_pack")
    _count(0)
    return out


# ---------------------------------------------------------------------------------------------------
# MLP
# ---------------------------------------------------------------------------------------------------


def mlp_fwd(rays_o, rays_d, z, packed, ins_num: int, mode: int, stash: bool = False):
    rays_o, rays_d, z = _f32(rays_o, "rays_o"), _f32(rays_d, "rays_d"), _f32(z, "z")
    n, s = z.shape
    raw = torch.empty((n, s, ins_num + 5), dtype=torch.float32, device=z.device)
    acts = None
    if stash:
        nbytes = _lib.lib().dmnerf_acts_bytes(n * s, ins_num, mode)
        if nbytes < 0:
            raise RuntimeError(f"mlp_fwd: mode {mode} keeps no activation stash (inference only); train in bf16 or fp32")
        acts = _aligned_bytes(nbytes, z.device)
    _timed_call(f"dmnerf_mlp_fwd[S={s},mode={mode}]", _lib.lib().dmnerf_mlp_fwd, _ptr(rays_o), _ptr(rays_d), _ptr(z),
                n, s, _ptr(packed), ins_num, mode, _ptr(raw), _ptr(acts), _stream())
    _count()
    return raw, acts


def mlp_density(rays_o, rays_d, z, packed, ins_num: int, mode: int) -> torch.Tensor:
    """sigma [N,S] = raw[..., 3] of ``mlp_fwd`` (same bits) from the trunk + density_linear alone (tensor-core modes)."""
    rays_o, rays_d, z = _f32(rays_o, "rays_o"), _f32(rays_d, "rays_d"), _f32(z, "z")
    n, s = z.shape
    sigma = torch.empty((n, s), dtype=torch.float32, device=z.device)
    _timed_call(f"dmnerf_mlp_density[S={s},mode={mode}]", _lib.lib().dmnerf_mlp_density, _ptr(rays_o), _ptr(rays_d), _ptr(z),
                n, s, _ptr(packed), ins_num, mode, _ptr(sigma), _stream())
    _count()
    return sigma


def weights_from_sigma(sigma, z, rays_d, channels: int) -> torch.Tensor:
    """The ``weights`` output of ``composite`` (render.py:9-14) from the density alone; ``channels`` = 5 + ins_num of the
    model picks the scan step the compositing kernel would use, so the bits are the same."""
    sigma, z, rays_d = _f32(sigma, "sigma"), _f32(z, "z_vals"), _f32(rays_d, "rays_d")
    n, s = sigma.shape
    w = torch.empty((n, s), dtype=torch.float32, device=sigma.device)
    _timed_call(f"dmnerf_weights_from_sigma[S={s}]", _lib.lib().dmnerf_weights_from_sigma, _ptr(sigma), _ptr(z), _ptr(rays_d),
                n, s, int(channels), _ptr(w), _stream())
    _count()
    return w


def coarse_tail(sigma, z_coarse, rays_d, channels: int, n_importance: int, u: Optional[torch.Tensor] = None,
                want_weights: bool = False):
    """``weights_from_sigma`` + ``fine_z`` in one launch (``dmnerf_coarse_tail``): z_fine [n, sc + n_importance] (and the
    coarse weights when asked).  ``z_coarse`` [n,sc], or one row [sc] / [1,sc] shared by every ray."""
    sigma, z_coarse, rays_d = _f32(sigma, "sigma"), _f32(z_coarse, "z_vals_coarse"), _f32(rays_d, "rays_d")
    n, sc = sigma.shape
    if z_coarse.numel() == sc:
        z_stride = 0
    else:
        assert z_coarse.shape == (n, sc), (z_coarse.shape, sigma.shape)
        z_stride = sc
    u, stride = _u_args(u, n, n_importance)
    z_fine = torch.empty((n, sc + n_importance), dtype=torch.float32, device=sigma.device)
    w = torch.empty((n, sc), dtype=torch.float32, device=sigma.device) if want_weights else None
    _timed_call("dmnerf_coarse_tail", _lib.lib().dmnerf_coarse_tail, _ptr(sigma), _ptr(z_coarse), z_stride, _ptr(rays_d),
                _ptr(u), stride, n, sc, int(channels), int(n_importance), _ptr(z_fine), _ptr(w), _stream())
    _count()
    return (z_fine, w) if want_weights else z_fine


def render_launches(count: int, chunk: int, mode: int, labels: bool = True) -> int:
    """Kernel launches one ``render_full`` call enqueues (for ``launches`` bookkeeping)."""
    if count <= 0:
        return 0
    passes = -(-count // chunk)
    per = 5 if mode != MODE_FP32 else 6
    return 1 + passes * (per + (1 if labels else 0))


def render_full(H: int, W: int, K, c2w, pix0: int, count: int, z_row: torch.Tensor, n_importance: int, packed_coarse,
                packed_fine, ins_num: int, mode: int, chunk: int, out, labels: bool = True,
                u_row: Optional[torch.Tensor] = None) -> None:
    """``dmnerf_render_full``: pixels [pix0, pix0+count) of the frame rendered into ``out`` (rgb [count,3], depth
    [count], ins [count,K], labels uint8 [count]) with every launch enqueued by the one C call."""
    if count <= 0:
        return
    dev = packed_fine.device
    z_row = _f32(z_row, "z_row").reshape(-1)
    sc = z_row.numel()
    if u_row is not None:
        u_row = _f32(u_row, "u_row").reshape(-1)
        assert u_row.numel() == n_importance
    kf = (C.c_float * 9)(*[float(v) for v in torch.as_tensor(K, dtype=torch.float32).cpu().reshape(-1)[:9]])
    cm = torch.as_tensor(c2w, dtype=torch.float32).cpu()[:3, :4].reshape(-1)
    cf = (C.c_float * 12)(*[float(v) for v in cm])
    need = int(_lib.lib().dmnerf_render_workspace_bytes(count, chunk, sc, int(n_importance), int(ins_num), int(mode)))
    if need < 0:
        raise RuntimeError(f"render_full: unsupported shape (n_samples {sc}, n_importance {n_importance}, ins_num {ins_num})")
    ws = torch.empty((need,), dtype=torch.uint8, device=dev)
    for k, shape, dt in (("rgb", (count, 3), torch.float32), ("depth", (count,), torch.float32),
                         ("ins", (count, ins_num), torch.float32)) + ((("labels", (count,), torch.uint8),) if labels else ()):
        t = out[k]
        if tuple(t.shape) != shape or t.dtype != dt or not t.is_contiguous() or t.device != dev:
            raise RuntimeError(f"render_full: out[{k!r}] must be a contiguous {dt} tensor of shape {shape} on {dev}")
    _timed_call("dmnerf_render_full", _lib.lib().dmnerf_render_full, int(H), int(W), kf, cf, int(pix0), int(count),
                _ptr(z_row), _ptr(u_row), sc, int(n_importance), _ptr(packed_coarse), _ptr(packed_fine), int(ins_num), int(mode),
                int(chunk), _ptr(ws), need, _ptr(out["rgb"]), _ptr(out["depth"]), _ptr(out["ins"]),
                _ptr(out["labels"]) if labels else C.c_void_p(0), _stream())
    passes = -(-count // chunk)
    _count(render_launches(count, -(-count // passes), mode, labels))


def mlp_fwd_embedded(x, packed, ins_num: int, mode: int, stash: bool = False):
    x = _f32(x, "x")
    if x.shape[-1] != 90:
        raise RuntimeError(f"DM_NeRF.forward: expected 63+27=90 input channels, got {x.shape[-1]}")
    lead = x.shape[:-1]
    x2 = x.reshape(-1, 90)
    m = x2.shape[0]
    raw = torch.empty((m, ins_num + 5), dtype=torch.float32, device=x.device)
    acts = None
    if stash:
        acts = _aligned_bytes(_lib.lib().dmnerf_acts_bytes(m, ins_num, mode), x.device)
    check(_lib.lib().dmnerf_mlp_fwd_embedded(_ptr(x2), m, _ptr(packed), ins_num, mode, _ptr(raw), _ptr(acts),
                                             _stream()), "dmnerf_mlp_fwd_embedded")
    _count()
    return raw.reshape(*lead, ins_num + 5), acts


def _aligned_bytes(nbytes: int, device) -> torch.Tensor:
    """uint8 buffer whose data pointer is 1024-byte aligned (bulk-copy / UMMA operand images)."""
    if nbytes < 0:
        raise RuntimeError(f"dm_nerf_b200: size query failed: {_lib.last_error()}")
    buf = torch.empty(nbytes + 1024, dtype=torch.uint8, device=device)
    off = (-buf.data_ptr()) % 1024
    return buf[off:off + nbytes]


def mlp_bwd(packed, acts, d_raw, n_points: int, ins_num: int, shapes: Sequence[torch.Size],
            mode: int = MODE_FP32) -> List[torch.Tensor]:
    d_raw = _f32(d_raw, "d_raw")
    # one zeroed buffer for the 30 gradients (the kernels accumulate into it): 1 fill instead of 30
    sizes = [int(torch.Size(s).numel()) for s in shapes]
    padded = [(n + 3) & ~3 for n in sizes]                                    # every gradient starts 16-byte aligned
    flat = torch.zeros(sum(padded), dtype=torch.float32, device=d_raw.device)
    grads = [g[:n].view(s) for g, n, s in zip(flat.split(padded), sizes, shapes)]
    ws_bytes = _lib.lib().dmnerf_bwd_workspace_bytes(n_points, ins_num, mode)
    ws = _aligned_bytes(ws_bytes, d_raw.device) if ws_bytes > 0 else None
    _timed_call(f"dmnerf_mlp_bwd[mode={mode}]", _lib.lib().dmnerf_mlp_bwd, _ptr(packed), _ptr(acts), _ptr(d_raw),
                n_points, ins_num, mode, _ptr_array(grads), _ptr(ws), _stream())
    _count(1 if mode == MODE_FP32 else 3)
    return grads


class _MlpFn(torch.autograd.Function):
    """raw = MLP(embed(o + d z), embed(d/|d|)) with parameter gradients (a12).  rays/z get no gradient
    (the reference detaches z_samples, render.py:68, and rays are data)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, z, x, ins_num, mode, packed, *params):
        if x is None:
            raw, acts = mlp_fwd(rays_o, rays_d, z, packed, ins_num, mode, stash=True)
            ctx.n_points = z.numel()
        else:
            raw, acts = mlp_fwd_embedded(x, packed, ins_num, mode, stash=True)
            ctx.n_points = x.numel() // 90
        ctx.save_for_backward(packed, acts)
        ctx.ins_num, ctx.mode = ins_num, mode
        ctx.shapes = [p.shape for p in params]
        return raw

    @staticmethod
    def backward(ctx, d_raw):
        packed, acts = ctx.saved_tensors
        grads = mlp_bwd(packed, acts, d_raw.reshape(-1, ctx.ins_num + 5), ctx.n_points, ctx.ins_num, ctx.shapes,
                        ctx.mode)
        return (None, None, None, None, None, None, None, *grads)


# ---------------------------------------------------------------------------------------------------
# compositing
# ---------------------------------------------------------------------------------------------------


def composite_fwd(raw, z, rays_d, flags: int, want_weights: bool = True, want_trans: bool = False):
    raw, z, rays_d = _f32(raw, "raw"), _f32(z, "z_vals"), _f32(rays_d, "rays_d")
    n, s, c = raw.shape
    dev = raw.device
    lout = c - 4 - (1 if flags & COMP_DROP_EMPTY else 0)
    rgb = torch.empty((n, 3), dtype=torch.float32, device=dev)
    depth = torch.empty((n,), dtype=torch.float32, device=dev)
    ins = torch.empty((n, lout), dtype=torch.float32, device=dev)
    weights = torch.empty((n, s), dtype=torch.float32, device=dev) if want_weights else None
    trans = torch.empty((n, s), dtype=torch.float32, device=dev) if want_trans else None
    _timed_call(f"dmnerf_composite_fwd[S={s}]", _lib.lib().dmnerf_composite_fwd, _ptr(raw), _ptr(z), _ptr(rays_d),
                n, s, c, flags, _ptr(rgb), _ptr(depth), _ptr(weights), _ptr(ins), _ptr(trans), _stream())
    _count()
    if want_trans:
        return rgb, weights, depth, ins, trans
    return rgb, weights, depth, ins


def hungarian(scores, n_valid):
    """Device-side scipy.optimize.linear_sum_assignment (evaluator.py:45-52).  scores [B, R, K] fp32 (rows beyond
    ``n_valid`` ignored), n_valid: int32 device scalar.  Returns int32 [B, K]: the matched column of each valid row,
    then the unmatched columns ascending.  Nothing is copied to the host."""
    scores = _f32(scores, "scores")
    b, r, k = scores.shape
    if n_valid.dtype != torch.int32 or not n_valid.is_cuda:
        raise RuntimeError("hungarian: n_valid must be an int32 CUDA tensor")
    col = torch.empty((b, k), dtype=torch.int32, device=scores.device)
    check(_lib.lib().dmnerf_hungarian(_ptr(scores), b, r * k, k, _ptr(n_valid), k, _ptr(col), _stream()), "dmnerf_hungarian")
    _count()
    return col


class _SegmentSum(torch.autograd.Function):
    """out[g, j] = sum_{n: slot[n] == g} x[n, j]  (``gt_one_hot.t() @ x`` without the dense product); backward is a gather."""

    @staticmethod
    def forward(ctx, x, slot, n_slots):
        x = _f32(x, "x")
        n, k = x.shape
        out = torch.empty((n_slots, k), dtype=torch.float32, device=x.device)
        check(_lib.lib().dmnerf_segment_sum(_ptr(x), _ptr(slot), n, k, n_slots, _ptr(out), _stream()), "dmnerf_segment_sum")
        _count()
        ctx.save_for_backward(slot)
        return out

    @staticmethod
    def backward(ctx, g):
        (slot,) = ctx.saved_tensors
        return g.index_select(0, slot), None, None


class _SegmentSum3(torch.autograd.Function):
    """(sum p, sum log(p + 1e-8), sum log(1 - p + 1e-8)) per ground-truth slot, one pass over p."""

    @staticmethod
    def forward(ctx, p, slot, perm, offs, n_slots):
        p = _f32(p, "pred_ins")
        n, k = p.shape
        out = torch.empty((3, n_slots, k), dtype=torch.float32, device=p.device)
        check(_lib.lib().dmnerf_segment_sum3(_ptr(p), _ptr(perm), _ptr(offs), n, k, n_slots, _ptr(out), _stream()),
              "dmnerf_segment_sum3")
        _count()
        ctx.save_for_backward(p, slot)
        return out[0], out[1], out[2]

    @staticmethod
    def backward(ctx, g_tp, g_lp, g_lq):
        p, slot = ctx.saved_tensors
        g = g_tp.index_select(0, slot) + g_lp.index_select(0, slot) / (p + 1e-8) - g_lq.index_select(0, slot) / (1 - p + 1e-8)
        return g, None, None, None, None


def segment_sum3(p, slot, perm, offs, n_slots: int):
    """(tp, sum log p, sum log(1-p)) label-wise for one prediction map (see ``dmnerf_segment_sum3``); differentiable in p.
    ``slot`` int64 [N] (the one-hot's column per ray), ``perm`` = its stable argsort, ``offs`` int64 [n_slots+1] = segment
    starts in ``perm`` -- all on the device, contiguous."""
    for name, t in (("slot", slot), ("perm", perm), ("offs", offs)):
        if t.dtype != torch.int64 or not t.is_cuda or not t.is_contiguous():
            raise RuntimeError(f"segment_sum3: {name} must be a contiguous int64 CUDA tensor")
    return _SegmentSum3.apply(p, slot, perm, offs, int(n_slots))


def ins_assign_loss_grad(p, slot, perm, offs, cnt, n_valid_i32, want_grad: bool = True):
    """The instance loss of ONE prediction map and its gradient as five launches (label-wise sums -> score matrices ->
    device Hungarian -> loss terms -> gradient) instead of ~130 tensor ops + autograd: evaluator.py:19-74 with a one-hot
    ground truth.  p [N,K] probabilities; slot/perm/offs/cnt/n_valid_i32 from ``ins_label_prep``.
    Returns (losses [4] = total, valid_ce, invalid_ce, valid_siou; grad [N,K] = d total / d p or None)."""
    p = _f32(p.detach(), "pred_ins")
    n, k = p.shape
    dev = p.device
    lib = _lib.lib()
    s3 = torch.empty((3, k, k), dtype=torch.float32, device=dev)
    check(lib.dmnerf_segment_sum3(_ptr(p), _ptr(perm), _ptr(offs), n, k, k, _ptr(s3), _stream()), "dmnerf_segment_sum3")
    mats = torch.empty((3, k, k), dtype=torch.float32, device=dev)               # scores | ce | siou
    colsum = torch.empty((2, k), dtype=torch.float32, device=dev)
    cnt = _f32(cnt, "gt_count")
    check(lib.dmnerf_assign_scores(_ptr(s3), _ptr(cnt), n, k, _ptr(mats[0]), _ptr(mats[1]), _ptr(mats[2]), _ptr(colsum),
                                   _stream()), "dmnerf_assign_scores")
    col = hungarian(mats[:1], n_valid_i32)[0]
    losses = torch.empty(4, dtype=torch.float32, device=dev)
    rowof = torch.empty(k, dtype=torch.int32, device=dev)
    grad = torch.empty((n, k), dtype=torch.float32, device=dev) if want_grad else None
    check(lib.dmnerf_assign_loss_grad(_ptr(p), _ptr(slot), _ptr(s3), _ptr(mats[1]), _ptr(mats[2]), _ptr(colsum), _ptr(cnt),
                                      _ptr(col), _ptr(n_valid_i32), n, k, _ptr(losses), _ptr(rowof), _ptr(grad), _stream()),
          "dmnerf_assign_loss_grad")
    _count(3)
    return losses, grad


def segment_sum(x, slot, n_slots: int):
    """Label-wise row sums of x [N,K] (evaluator.py:55-63 with a one-hot ground truth): [n_slots, K].  ``slot`` int64 [N]
    on the device.  Differentiable in x."""
    if slot.dtype != torch.int64 or not slot.is_cuda or not slot.is_contiguous():
        raise RuntimeError("segment_sum: slot must be a contiguous int64 CUDA tensor")
    return _SegmentSum.apply(x, slot, int(n_slots))


def hungarian_infeasible() -> int:
    """Launches of ``hungarian`` since the last call whose score matrix had no assignment (a row of NaN / inf; scipy
    raises ValueError there).  Synchronises the current stream -- call it outside captured graphs."""
    cnt = C.c_int(0)
    check(_lib.lib().dmnerf_hungarian_infeasible(C.byref(cnt), _stream()), "dmnerf_hungarian_infeasible")
    return int(cnt.value)


def composite_bwd(raw, z, rays_d, flags, ins_detach, d_rgb, d_depth, d_weights, d_ins, fwd=None):
    """``fwd`` = (weights, trans, ins) kept from the forward -> one reverse pass instead of recompute + reverse."""
    n, s, c = raw.shape
    d_raw = torch.empty_like(raw)
    opt = lambda t: None if t is None else _f32(t, "grad")
    d_rgb, d_depth, d_weights, d_ins = opt(d_rgb), opt(d_depth), opt(d_weights), opt(d_ins)
    fw, ft, fi = fwd if fwd is not None else (None, None, None)
    check(_lib.lib().dmnerf_composite_bwd(_ptr(raw), _ptr(z), _ptr(rays_d), n, s, c, flags, int(ins_detach),
                                          _ptr(d_rgb), _ptr(d_depth), _ptr(d_weights), _ptr(d_ins), _ptr(d_raw),
                                          _ptr(fw), _ptr(ft), _ptr(fi), _stream()), "dmnerf_composite_bwd")
    _count()
    return d_raw


class _CompositeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw, z, rays_d, flags, ins_detach):
        raw, z, rays_d = _f32(raw, "raw"), _f32(z, "z_vals"), _f32(rays_d, "rays_d")
        rgb, weights, depth, ins, trans = composite_fwd(raw, z, rays_d, flags, want_trans=True)
        ctx.save_for_backward(raw, z, rays_d, weights, trans, ins)
        ctx.flags, ctx.ins_detach = flags, ins_detach
        return rgb, weights, depth, ins

    @staticmethod
    def backward(ctx, d_rgb, d_weights, d_depth, d_ins):
        raw, z, rays_d, weights, trans, ins = ctx.saved_tensors
        d_raw = composite_bwd(raw, z, rays_d, ctx.flags, ctx.ins_detach, d_rgb, d_depth, d_weights, d_ins,
                              fwd=(weights, trans, ins))
        return d_raw, None, None, None, None


def composite(raw, z, rays_d, flags: int, ins_detach: bool):
    if torch.is_grad_enabled() and isinstance(raw, torch.Tensor) and raw.requires_grad:
        return _CompositeFn.apply(raw, z, rays_d, flags, ins_detach)
    return composite_fwd(raw, z, rays_d, flags)


# ---------------------------------------------------------------------------------------------------
# sampler
# ---------------------------------------------------------------------------------------------------


def _u_args(u: Optional[torch.Tensor], n: int, ns: int):
    if u is None:
        return None, 0
    u = _f32(u, "u")
    if u.dim() == 1:
        assert u.shape[0] == ns
        return u, 0
    assert tuple(u.shape) == (n, ns), (u.shape, n, ns)
    return u, ns


def sample_pdf(bins, weights, n_samples: int, u: Optional[torch.Tensor], want_inds: bool = False,
               cdf: Optional[torch.Tensor] = None):
    bins = _f32(bins, "bins")
    n, nb = bins.shape
    u, stride = _u_args(u, n, n_samples)
    z = torch.empty((n, n_samples), dtype=torch.float32, device=bins.device)
    inds = torch.empty((n, n_samples), dtype=torch.int64, device=bins.device) if want_inds else None
    if cdf is not None:
        cdf = _f32(cdf, "cdf")
        check(_lib.lib().dmnerf_sample_cdf(_ptr(bins), _ptr(cdf), _ptr(u), stride, n, nb, n_samples, _ptr(z),
                                           _ptr(inds), _stream()), "dmnerf_sample_cdf")
    else:
        weights = _f32(weights, "weights")
        assert weights.shape == (n, nb - 1), (weights.shape, bins.shape)
        check(_lib.lib().dmnerf_sample_pdf(_ptr(bins), _ptr(weights), _ptr(u), stride, n, nb, n_samples, _ptr(z),
                                           _ptr(inds), _stream()), "dmnerf_sample_pdf")
    _count()
    return (z, inds) if want_inds else z


def fine_z(z_coarse, weights, n_importance: int, u: Optional[torch.Tensor], want_samples: bool = False):
    z_coarse, weights = _f32(z_coarse, "z_vals_coarse"), _f32(weights, "weights")
    n, sc = z_coarse.shape
    u, stride = _u_args(u, n, n_importance)
    z_fine = torch.empty((n, sc + n_importance), dtype=torch.float32, device=z_coarse.device)
    zs = torch.empty((n, n_importance), dtype=torch.float32, device=z_coarse.device) if want_samples else None
    _timed_call("dmnerf_fine_z", _lib.lib().dmnerf_fine_z, _ptr(z_coarse), _ptr(weights), _ptr(u), stride, n, sc,
                n_importance, _ptr(z_fine), _ptr(zs), _stream())
    _count()
    return (z_fine, zs) if want_samples else z_fine


def sort_cat(parts: Sequence[torch.Tensor]) -> torch.Tensor:
    parts = [_f32(p, "part") for p in parts]
    n = parts[0].shape[0]
    lens = (C.c_int * len(parts))(*[p.shape[1] for p in parts])
    out = torch.empty((n, sum(p.shape[1] for p in parts)), dtype=torch.float32, device=parts[0].device)
    check(_lib.lib().dmnerf_sort_cat(_ptr_array(parts), lens, len(parts), n, _ptr(out), _stream()), "dmnerf_sort_cat")
    _count()
    return out


def jitter_z(z, t_rand):
    z, t_rand = _f32(z, "z_vals"), _f32(t_rand, "t_rand")
    out = torch.empty_like(z)
    check(_lib.lib().dmnerf_jitter_z(_ptr(z), _ptr(t_rand), z.shape[0], z.shape[1], _ptr(out), _stream()),
          "dmnerf_jitter_z")
    _count()
    return out


# ---------------------------------------------------------------------------------------------------
# penalizer
# ---------------------------------------------------------------------------------------------------


def penalizer_fwd(raw, z, depth, rays_d, tolerance: float, deta_w: float, flags: int):
    n, s, c = raw.shape
    partials = torch.empty(4, dtype=torch.float32, device=raw.device)
    loss = torch.empty(1, dtype=torch.float32, device=raw.device)
    check(_lib.lib().dmnerf_penalizer_fwd(_ptr(raw), _ptr(z), _ptr(depth), _ptr(rays_d), n, s, c, tolerance, deta_w,
                                          flags, _ptr(partials), _ptr(loss), _stream()), "dmnerf_penalizer_fwd")
    _count(2)
    return loss, partials


def penalizer_bwd(raw, z, depth, rays_d, tolerance, deta_w, flags, partials, d_loss, d_raw):
    n, s, c = raw.shape
    check(_lib.lib().dmnerf_penalizer_bwd(_ptr(raw), _ptr(z), _ptr(depth), _ptr(rays_d), n, s, c, tolerance, deta_w,
                                          flags, _ptr(partials), _ptr(d_loss), _ptr(d_raw), _stream()),
          "dmnerf_penalizer_bwd")
    _count()
    return d_raw


class _PenalizerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw, z, depth, rays_d, tolerance, deta_w):
        raw, z, depth, rays_d = _f32(raw, "raw"), _f32(z, "z_vals"), _f32(depth, "depth"), _f32(rays_d, "rays_d")
        ctx.flags = 0
        loss, partials = penalizer_fwd(raw, z, depth, rays_d, tolerance, deta_w, 0)
        ctx.save_for_backward(raw, z, depth, rays_d, partials)
        ctx.tol, ctx.dw = tolerance, deta_w
        return loss

    @staticmethod
    def backward(ctx, d_loss):
        raw, z, depth, rays_d, partials = ctx.saved_tensors
        d_loss = _f32(d_loss, "d_loss").reshape(1)
        d_raw = penalizer_bwd(raw, z, depth, rays_d, ctx.tol, ctx.dw, ctx.flags, partials, d_loss, torch.empty_like(raw))
        return d_raw, None, None, None, None, None


def penalizer(raw, z, depth, rays_d, tolerance: float, deta_w: float) -> torch.Tensor:
    return _PenalizerFn.apply(raw, z, depth.reshape(-1).detach(), rays_d, float(tolerance), float(deta_w))


class _CompositePenalizedFn(torch.autograd.Function):
    """``render_train`` (render.py:6-28) and ``ins_penalizer`` (penalizer.py:58-62, on the composited depth, detached)
    of the same ``raw`` as ONE autograd node: the backward writes the compositing gradient and the penalizer adds its
    own into it in place (DMNERF_PEN_ACCUMULATE) -- no zero-filled second gradient, no [N,S,C] gradient-sum pass."""

    @staticmethod
    def forward(ctx, raw, z, rays_d, flags, ins_detach, tolerance, deta_w, after_composite=None):
        raw, z, rays_d = _f32(raw, "raw"), _f32(z, "z_vals"), _f32(rays_d, "rays_d")
        rgb, weights, depth, ins, trans = composite_fwd(raw, z, rays_d, flags, want_trans=True)
        if after_composite is not None:
            after_composite(ins)          # the composited map exists; the penalizer kernel has not been enqueued yet
        ctx.pen_flags = 0
        loss, partials = penalizer_fwd(raw, z, depth, rays_d, tolerance, deta_w, 0)
        ctx.save_for_backward(raw, z, rays_d, depth, partials, weights, trans, ins)
        ctx.flags, ctx.ins_detach, ctx.tol, ctx.dw = flags, ins_detach, tolerance, deta_w
        return rgb, weights, depth, ins, loss

    @staticmethod
    def backward(ctx, d_rgb, d_weights, d_depth, d_ins, d_loss):
        raw, z, rays_d, depth, partials, weights, trans, ins = ctx.saved_tensors
        d_raw = composite_bwd(raw, z, rays_d, ctx.flags, ctx.ins_detach, d_rgb, d_depth, d_weights, d_ins,
                              fwd=(weights, trans, ins))
        if d_loss is not None:
            d_loss = _f32(d_loss, "d_loss").reshape(1)
            penalizer_bwd(raw, z, depth, rays_d, ctx.tol, ctx.dw, ctx.pen_flags | _lib.PEN_ACCUMULATE, partials, d_loss, d_raw)
        return d_raw, None, None, None, None, None, None, None


def composite_penalized(raw, z, rays_d, flags: int, ins_detach: bool, tolerance: float, deta_w: float, after_composite=None):
    """(rgb, weights, depth, ins, emptiness_loss); see _CompositePenalizedFn.  ``after_composite(ins)`` is called between
    the compositing launch and the penalizer launch (grad mode is off there)."""
    return _CompositePenalizedFn.apply(raw, z, rays_d, flags, ins_detach, float(tolerance), float(deta_w), after_composite)


# ---------------------------------------------------------------------------------------------------
# exchanger / rays / labels
# ---------------------------------------------------------------------------------------------------


def exchange(ori_raw, tar_raw, ori_acc, tar_acc, move_label: int, ori_label: Optional[torch.Tensor],
             want_ori_label: bool = True, want_tar_label: bool = True):
    """One move label, in place on ori_raw.  Returns (ori_label, tar_label) int64 [n,s]; a label tensor the caller
    does not want (``want_*_label=False``; ori only on the first label) is neither allocated nor written."""
    if not (ori_raw.is_cuda and ori_raw.dtype == torch.float32 and ori_raw.is_contiguous()):
        raise RuntimeError("exchanger: ori_raw must be a contiguous fp32 CUDA tensor (it is edited in place)")
    tar_raw, ori_acc, tar_acc = _f32(tar_raw, "tar_raw"), _f32(ori_acc, "ori_raw_pred"), _f32(tar_acc, "tar_raw_pred")
    n, s, c = ori_raw.shape
    first = ori_label is None
    if first and want_ori_label:
        ori_label = torch.empty((n, s), dtype=torch.int64, device=ori_raw.device)
    tar_label = torch.empty((n, s), dtype=torch.int64, device=ori_raw.device) if want_tar_label else None
    _timed_call(f"dmnerf_exchange[S={s}]", _lib.lib().dmnerf_exchange, _ptr(ori_raw), _ptr(tar_raw), _ptr(ori_acc),
                _ptr(tar_acc), int(move_label), int(first), n, s, c, _ptr(ori_label), _ptr(tar_label), _stream())
    _count()
    return ori_label, tar_label


def embed(x, n_freqs: int) -> torch.Tensor:
    x = _f32(x, "inputs")
    lead = x.shape[:-1]
    x2 = x.reshape(-1, 3)
    out = torch.empty((x2.shape[0], 3 + 6 * n_freqs), dtype=torch.float32, device=x.device)
    check(_lib.lib().dmnerf_embed(_ptr(x2), x2.shape[0], n_freqs, _ptr(out), _stream()), "dmnerf_embed")
    _count()
    return out.reshape(*lead, 3 + 6 * n_freqs)


def rays_k(H: int, W: int, K, c2w, pix0: int, n: int, device, pix_idx: Optional[torch.Tensor] = None
           ) -> Tuple[torch.Tensor, torch.Tensor]:
    kf = (C.c_float * 9)(*[float(v) for v in torch.as_tensor(K, dtype=torch.float32).cpu().reshape(-1)[:9]])
    cm = torch.as_tensor(c2w, dtype=torch.float32).cpu()[:3, :4].reshape(-1)
    cf = (C.c_float * 12)(*[float(v) for v in cm])
    ro = torch.empty((n, 3), dtype=torch.float32, device=device)
    rd = torch.empty((n, 3), dtype=torch.float32, device=device)
    if pix_idx is not None:
        pix_idx = pix_idx.to(device=device, dtype=torch.int64).contiguous()
    check(_lib.lib().dmnerf_rays_k(H, W, kf, cf, pix0, _ptr(pix_idx), n, _ptr(ro), _ptr(rd), _stream()),
          "dmnerf_rays_k")
    _count()
    return ro, rd


def rays_k_dev(H: int, W: int, K, c2w, n: int, pix_idx: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``rays_k`` with the intrinsics and the pose as CUDA tensors (no host copy, nothing view-specific in the launch)."""
    K = _f32(K, "K")
    c2w = _f32(c2w, "c2w")
    if K.numel() < 9 or c2w.numel() < 12 or c2w.shape[-1] != 4:
        raise RuntimeError("rays_k_dev: K must be [3,3] and c2w [>=3,4]")
    dev = K.device
    ro = torch.empty((n, 3), dtype=torch.float32, device=dev)
    rd = torch.empty((n, 3), dtype=torch.float32, device=dev)
    pix_idx = pix_idx.to(device=dev, dtype=torch.int64).contiguous()
    check(_lib.lib().dmnerf_rays_k_dev(H, W, _ptr(K), _ptr(c2w), 0, _ptr(pix_idx), n, _ptr(ro), _ptr(rd), _stream()),
          "dmnerf_rays_k_dev")
    _count()
    return ro, rd


def draw_pixels(n: int, total: int, step: torch.Tensor, seed: int) -> torch.Tensor:
    """n distinct pixel indices in [0,total) for the training step whose number sits in the device int32 ``step``."""
    if step.dtype != torch.int32 or not step.is_cuda:
        raise RuntimeError("draw_pixels: step must be an int32 CUDA tensor")
    idx = torch.empty((n,), dtype=torch.int64, device=step.device)
    check(_lib.lib().dmnerf_draw_pixels(_ptr(idx), n, total, _ptr(step), C.c_uint64(seed & 0xFFFFFFFFFFFFFFFF), _stream()),
          "dmnerf_draw_pixels")
    _count()
    return idx


def label_argmax(ins) -> torch.Tensor:
    ins = _f32(ins, "ins")
    n, k = ins.shape
    out = torch.empty((n,), dtype=torch.uint8, device=ins.device)
    check(_lib.lib().dmnerf_label_argmax(_ptr(ins), n, k, _ptr(out), _stream()), "dmnerf_label_argmax")
    _count()
    return out
