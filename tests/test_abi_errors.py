"""The C-ABI's error paths on a real device (include/dmnerf_b200.h: int return codes + dmnerf_last_error), and the
edge cases round 1's review found unguarded: K = 1 in the penalizer's index split, a score matrix without an
assignment in the device Hungarian.  A failed call must leave the library and the stream usable."""
import ctypes as C

import numpy as np
import pytest
import torch

import dmnerf_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _codes():
    """DMNERF_EINVAL / DMNERF_EUNSUPPORTED as the header defines them."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "dmnerf_b200.h")).read()
    got = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(DMNERF_E\w+)\s+(-?\d+)", text)}
    return got


def test_error_codes_and_messages_then_stream_still_works():
    from dm_nerf_b200 import _lib, ops
    lib = _lib.lib()
    codes = _codes()
    einval, eunsup = codes["DMNERF_EINVAL"], codes["DMNERF_EUNSUPPORTED"]
    n, s, k = 8, 64, 32
    c = k + 5
    raw = torch.randn(n, s, c, device=DEV)
    z = torch.sort(torch.rand(n, s, device=DEV) * 5 + 2, -1)[0]
    rd = torch.randn(n, 3, device=DEV)
    rgb, depth, w, ins = (torch.empty(n, 3, device=DEV), torch.empty(n, device=DEV), torch.empty(n, s, device=DEV),
                          torch.empty(n, k, device=DEV))

    def comp(raw_t, s_, c_):
        return lib.dmnerf_composite_fwd(_p(raw_t), _p(z), _p(rd), n, s_, c_, 1, _p(rgb), _p(depth), _p(w), _p(ins), _p(None), _st())

    # null pointer -> EINVAL with a message
    rc = comp(None, s, c)
    assert rc == einval and "null" in _lib.last_error().lower(), (rc, _lib.last_error())
    # n_samples above the supported range -> EUNSUPPORTED
    rc = comp(raw, 1025, c)
    assert rc == eunsup and _lib.last_error(), (rc, _lib.last_error())
    # ins_num above the supported range (C = ins_num + 5) -> EUNSUPPORTED, also through the size queries
    rc = comp(raw, s, 124 + 5)
    assert rc == eunsup and "channels" in _lib.last_error().lower(), (rc, _lib.last_error())
    assert lib.dmnerf_packed_bytes(124, 0) == -1 and lib.dmnerf_packed_bytes(0, 1) == -1
    assert lib.dmnerf_acts_bytes(1024, 124, 1) < 0
    # misaligned packed image in the tensor-core mode -> EINVAL
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF
    net = DM_NeRF(8, 256, 63, 27, [4], k)
    net.load_state_dict(orc.init_params(k, 1))
    net = net.to(DEV)
    packed = net.packed(_lib.MODE_BF16)
    shifted = torch.empty(packed.numel() + 16, dtype=torch.uint8, device=DEV)
    shifted[4:4 + packed.numel()].copy_(packed)
    ro = torch.zeros(n, 3, device=DEV)
    out = torch.empty(n, s, c, device=DEV)
    rc = lib.dmnerf_mlp_fwd(_p(ro), _p(rd), _p(z), n, s, C.c_void_p(shifted.data_ptr() + 4), k, _lib.MODE_BF16, _p(out),
                            _p(None), _st())
    assert rc == einval and "align" in _lib.last_error().lower(), (rc, _lib.last_error())
    # unknown mode -> error, not a crash
    rc = lib.dmnerf_mlp_fwd(_p(ro), _p(rd), _p(z), n, s, _p(packed), k, 7, _p(out), _p(None), _st())
    assert rc != 0 and _lib.last_error()
    # the Python layer turns codes into RuntimeError carrying code and text
    with pytest.raises(RuntimeError, match="code"):
        _lib.check(comp(None, s, c), "dmnerf_composite_fwd")
    # ... and after all those refusals the same stream still computes correctly
    assert comp(raw, s, c) == 0
    ref = orc.render_train(raw.cpu(), z.cpu(), rd.cpu())
    torch.cuda.synchronize()
    assert float((rgb.cpu() - ref[0]).abs().max()) <= 2e-5
    got = ops.mlp_fwd(ro, rd, z, packed, k, _lib.MODE_BF16)[0]
    assert torch.isfinite(got).all()


def test_penalizer_single_instance():
    """K = 1 (C = 6): the (listed sample, channel) index split degenerates to one channel per sample
    (penalizer.cu; 0xFFFFFFFF / 1 + 1 wraps to 0)."""
    from dm_nerf_b200 import ops
    gen = torch.Generator().manual_seed(9)
    n, s, k = 13, 96, 1
    raw = torch.randn(n, s, k + 5, generator=gen)
    z = torch.sort(torch.rand(n, s, generator=gen) * 6 + 2, -1)[0]
    depth = 2.5 + 5 * torch.rand(n, generator=gen)
    rd = torch.randn(n, 3, generator=gen)
    a = raw.clone().requires_grad_(True)
    ref = orc.ins_penalizer(a, z, depth, rd, 0.05, 0.05)
    ref.sum().backward()
    b = raw.to(DEV).requires_grad_(True)
    got = ops.penalizer(b, z.to(DEV), depth.to(DEV), rd.to(DEV), 0.05, 0.05)
    got.sum().backward()
    assert abs(float(got) - float(ref)) <= 1e-5 * max(1.0, abs(float(ref)))
    gmax = float(a.grad.abs().max())
    assert float((b.grad.cpu() - a.grad).abs().max()) <= 1e-7 + 1e-4 * gmax
    assert torch.all(b.grad[..., :4] == 0)


def test_hungarian_without_assignment_is_defined():
    """A NaN / inf row (a diverged step): scipy raises ValueError; the kernel writes the identity permutation -- every
    output slot defined and in range -- and the launch is counted."""
    from dm_nerf_b200 import ops
    k = 16
    gen = torch.Generator().manual_seed(2)
    scores = torch.rand(3, k, k, generator=gen)
    scores[1, 4, :] = float("nan")
    scores[2, 0, :] = float("inf")
    ops.hungarian_infeasible()                                                # reset
    col = ops.hungarian(scores.to(DEV), torch.tensor(k, dtype=torch.int32, device=DEV)).cpu().numpy()
    assert ops.hungarian_infeasible() == 2 and ops.hungarian_infeasible() == 0
    from scipy.optimize import linear_sum_assignment
    assert col[0].tolist() == linear_sum_assignment(scores[0].double().numpy())[1].tolist()
    for b in (1, 2):
        assert col[b].tolist() == list(range(k))
    assert ((col >= 0) & (col < k)).all()


def test_inference_only_modes_refuse_training_calls():
    """DMNERF_MODE_F16 / DMNERF_MODE_TC32 are inference chains: size queries for a stash or a backward workspace say -1,
    a forward that is handed an activation stash and a backward in those modes return DMNERF_EUNSUPPORTED with a message,
    an unknown mode is refused everywhere, and the stream keeps working afterwards.  exchange: ori_label may be NULL
    only on the first label."""
    from dm_nerf_b200 import _lib
    lib = _lib.lib()
    codes = _codes()
    einval, eunsup = codes["DMNERF_EINVAL"], codes["DMNERF_EUNSUPPORTED"]
    k, n, s = 8, 4, 16
    sd = orc.init_params(k, 3)
    params = [sd[name].to(DEV).contiguous() for name in __import__("dm_nerf_b200.ops", fromlist=["PARAM_ORDER"]).PARAM_ORDER]
    arr = (C.c_void_p * len(params))(*[p.data_ptr() for p in params])
    o, d = torch.rand(n, 3, device=DEV), torch.randn(n, 3, device=DEV)
    z = torch.sort(torch.rand(n, s, device=DEV) * 5 + 2, -1)[0]
    raw = torch.empty(n, s, k + 5, device=DEV)
    scratch = torch.empty(1 << 20, dtype=torch.uint8, device=DEV)
    for mode in (_lib.MODE_F16, _lib.MODE_TC32):
        nbytes = lib.dmnerf_packed_bytes(k, mode)
        assert nbytes > 0
        assert lib.dmnerf_acts_bytes(n * s, k, mode) == -1 and lib.dmnerf_bwd_workspace_bytes(n * s, k, mode) == -1
        packed = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
        assert lib.dmnerf_weights_pack(arr, k, mode, _p(packed), _st()) == 0
        rc = lib.dmnerf_mlp_fwd(_p(o), _p(d), _p(z), n, s, _p(packed), k, mode, _p(raw), _p(scratch), _st())
        assert rc == eunsup and "inference" in _lib.last_error().lower(), (rc, _lib.last_error())
        rc = lib.dmnerf_mlp_bwd(_p(packed), _p(scratch), _p(raw), n * s, k, mode, arr, _p(None), _st())
        assert rc == eunsup and _lib.last_error(), (rc, _lib.last_error())
        assert lib.dmnerf_mlp_fwd(_p(o), _p(d), _p(z), n, s, _p(packed), k, mode, _p(raw), _p(None), _st()) == 0
        torch.cuda.synchronize()
        assert bool(torch.isfinite(raw).all())
    assert lib.dmnerf_packed_bytes(k, 7) == -1
    rc = lib.dmnerf_mlp_fwd(_p(o), _p(d), _p(z), n, s, _p(scratch), k, 7, _p(raw), _p(None), _st())
    assert rc == eunsup and "mode" in _lib.last_error().lower()
    # exchange without label maps: tar_label NULL is fine, ori_label NULL only with first_label != 0
    c = k + 5
    ori, tar = torch.randn(n, s, c, device=DEV), torch.randn(n, s, c, device=DEV)
    acc = torch.rand(n, k + 1, device=DEV)
    assert lib.dmnerf_exchange(_p(ori), _p(tar), _p(acc), _p(acc), 1, 1, n, s, c, _p(None), _p(None), _st()) == 0
    rc = lib.dmnerf_exchange(_p(ori), _p(tar), _p(acc), _p(acc), 1, 0, n, s, c, _p(None), _p(None), _st())
    assert rc == einval and "ori_label" in _lib.last_error(), (rc, _lib.last_error())
    torch.cuda.synchronize()


def test_render_full_refuses_bad_arguments():
    """dmnerf_render_full / dmnerf_coarse_tail (the calls a frame driver makes once per pixel range): short or misaligned
    workspace, pixels outside the frame, null buffers and unsupported shapes come back as codes with a message, and the
    call after them renders."""
    from dm_nerf_b200 import _lib, ops, synth
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF
    lib = _lib.lib()
    codes = _codes()
    einval, eunsup = codes["DMNERF_EINVAL"], codes["DMNERF_EUNSUPPORTED"]
    k, H, W, sc, n_imp = 32, 8, 16, 64, 128
    net = DM_NeRF(8, 256, 63, 27, [4], k)
    net.load_state_dict(orc.init_params(k, 1))
    net = net.to(DEV)
    mode = _lib.MODE_F16
    packed = net.packed(mode)
    Kmat, c2w = synth.dmsr_camera(H, W, 0)
    kf = (C.c_float * 9)(*[float(v) for v in torch.as_tensor(Kmat, dtype=torch.float32).reshape(-1)[:9]])
    cf = (C.c_float * 12)(*[float(v) for v in torch.as_tensor(c2w, dtype=torch.float32)[:3, :4].reshape(-1)])
    count = H * W
    z_row = (4.0 + torch.linspace(0., 1., sc) * 11.0).to(DEV)
    need = lib.dmnerf_render_workspace_bytes(count, 4096, sc, n_imp, k, mode)
    assert need > count * (sc + n_imp) * (k + 5) * 4
    assert lib.dmnerf_render_workspace_bytes(count, 4096, sc, n_imp, 124, mode) == -1
    assert lib.dmnerf_render_workspace_bytes(0, 4096, sc, n_imp, k, mode) == 0
    ws = torch.empty(need + 256, dtype=torch.uint8, device=DEV)
    rgb, depth, ins = torch.empty(count, 3, device=DEV), torch.empty(count, device=DEV), torch.empty(count, k, device=DEV)
    lab = torch.empty(count, dtype=torch.uint8, device=DEV)

    def run(pix0=0, cnt=count, ws_ptr=None, ws_bytes=need, rgb_t=rgb, n_imp_=n_imp, mode_=mode):
        ptr = C.c_void_p(ws.data_ptr() if ws_ptr is None else ws_ptr)
        return lib.dmnerf_render_full(H, W, kf, cf, pix0, cnt, _p(z_row), _p(None), sc, n_imp_, _p(packed), _p(packed), k, mode_,
                                      4096, ptr, ws_bytes, _p(rgb_t), _p(depth), _p(ins), _p(lab), _st())

    assert run(ws_bytes=need - 1) == einval and "workspace" in _lib.last_error().lower()
    assert run(ws_ptr=ws.data_ptr() + 4) == einval and "align" in _lib.last_error().lower()
    assert run(pix0=1) == einval and "frame" in _lib.last_error().lower()
    assert run(rgb_t=None) == einval and "null" in _lib.last_error().lower()
    assert run(n_imp_=2000) == eunsup and _lib.last_error()
    assert run(mode_=9) == eunsup and "mode" in _lib.last_error().lower()
    assert run(cnt=0) == 0
    sigma = torch.rand(4, sc, device=DEV)
    rd = torch.randn(4, 3, device=DEV)
    zf = torch.empty(4, sc + n_imp, device=DEV)
    rc = lib.dmnerf_coarse_tail(_p(sigma), _p(z_row), 3, _p(rd), _p(None), 0, 4, sc, k + 5, n_imp, _p(zf), _p(None), _st())
    assert rc == einval and "z_stride" in _lib.last_error()
    rc = lib.dmnerf_coarse_tail(_p(None), _p(z_row), 0, _p(rd), _p(None), 0, 4, sc, k + 5, n_imp, _p(zf), _p(None), _st())
    assert rc == einval and "null" in _lib.last_error().lower()
    # ... and the next call renders
    assert run() == 0
    torch.cuda.synchronize()
    assert torch.isfinite(rgb).all() and torch.isfinite(ins).all() and int(lab.max()) < k
