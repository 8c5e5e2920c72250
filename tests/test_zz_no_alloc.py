"""dmnerf_mlp_bwd_bf16_noalloc (include/dmnerf_b200.h): the bf16 MLP backward with its workspace in a static __device__
array inside the library -- no allocation in the call.  Called through the C-ABI and checked against the existing entry
point (dmnerf_mlp_bwd with a caller-provided workspace) on two calls in a row; the second call finds the first call's
contents in the shared scratch and must not be affected by them.

Collected last on purpose (test_zz_*).  Passes alone in a fresh process and after
the other tests (it shares no state with them but the loaded library)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr_array(tensors):
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def _setup(k, seed):
    from dm_nerf_b200 import _lib, ops, synth
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF
    m = DM_NeRF(8, 256, 63, 27, [4], k)
    m.load_state_dict(synth.init_state_dict(k, seed))
    m = m.to(DEV)
    return m, m.packed(_lib.MODE_BF16), [p.shape for p in m._ordered_params()], ops, _lib


def _forward(ops, _lib, packed, k, n_rays, s, seed):
    gen = torch.Generator().manual_seed(seed)
    o = ((torch.rand(n_rays, 3, generator=gen) - 0.5) * 2).to(DEV)
    d = torch.randn(n_rays, 3, generator=gen).to(DEV)
    z = torch.sort(torch.rand(n_rays, s, generator=gen) * 11 + 4, -1)[0].to(DEV)
    raw, acts = ops.mlp_fwd(o, d, z, packed, k, _lib.MODE_BF16, stash=True)
    d_raw = torch.randn(n_rays * s, k + 5, generator=gen).to(DEV)
    return acts, d_raw


def _zero_grads(shapes):
    return [torch.zeros(tuple(sh), dtype=torch.float32, device=DEV) for sh in shapes]


def _existing(lib, _lib, ops, packed, acts, d_raw, n_points, k, shapes):
    grads = _zero_grads(shapes)
    ws = ops._aligned_bytes(lib.dmnerf_bwd_workspace_bytes(n_points, k, _lib.MODE_BF16), DEV)
    rc = lib.dmnerf_mlp_bwd(_p(packed), _p(acts), _p(d_raw), n_points, k, _lib.MODE_BF16, _ptr_array(grads), _p(ws), _st())
    assert rc == 0, _lib.last_error()
    return grads


def _noalloc(lib, _lib, packed, acts, d_raw, n_points, k, shapes):
    grads = _zero_grads(shapes)                       # inputs and outputs stay the caller's; only the scratch is static
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    rc = lib.dmnerf_mlp_bwd_bf16_noalloc(_p(packed), _p(acts), _p(d_raw), n_points, k, _ptr_array(grads), _st())
    assert rc == 0, _lib.last_error()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before      # nothing went through torch's allocator
    return grads


def _close(got, ref):
    # the tolerance tests/test_gpu_parity.py uses between two runs of the same wgrad kernels (only the order of their
    # floating-point reductions differs): 2e-4 of the largest gradient entry
    for i, (g, r) in enumerate(zip(got, ref)):
        tol = 1e-7 + 2e-4 * float(r.abs().max())
        err = float((g - r).abs().max())
        assert err <= tol, f"param {i}: max abs diff {err} > {tol}"
    assert any(float(r.abs().max()) > 0 for r in ref)


def test_noalloc_backward_matches_existing_entry_point_twice():
    k = 32
    m, packed, shapes, ops, _lib = _setup(k, 3)
    lib = _lib.lib()
    max_pts = int(lib.dmnerf_noalloc_max_points())
    assert max_pts == 2048
    # call 1: the maximum size (32 rays x 64 samples); call 2: a smaller, ragged size (7 rays x 77 samples = 539 points: a
    # partial 256-point pair) right after it, so the scratch still holds call 1's stash and gradient image
    for n_rays, s, seed in ((32, 64, 11), (7, 77, 12)):
        n_points = n_rays * s
        assert n_points <= max_pts
        acts, d_raw = _forward(ops, _lib, packed, k, n_rays, s, seed)
        ref = _existing(lib, _lib, ops, packed, acts, d_raw, n_points, k, shapes)
        got = _noalloc(lib, _lib, packed, acts, d_raw, n_points, k, shapes)
        torch.cuda.synchronize()
        _close(got, ref)
    # and once more at the first size with another ins_num: same library-wide scratch, different packed layout
    k2 = 93
    m2, packed2, shapes2, _, _ = _setup(k2, 4)
    acts, d_raw = _forward(ops, _lib, packed2, k2, 16, 128, 13)
    ref = _existing(lib, _lib, ops, packed2, acts, d_raw, 2048, k2, shapes2)
    got = _noalloc(lib, _lib, packed2, acts, d_raw, 2048, k2, shapes2)
    torch.cuda.synchronize()
    _close(got, ref)


def test_noalloc_backward_rejects_sizes_beyond_the_static_scratch():
    k = 32
    m, packed, shapes, ops, _lib = _setup(k, 3)
    lib = _lib.lib()
    n_rays, s = 33, 64                                               # 2112 points > 2048
    acts, d_raw = _forward(ops, _lib, packed, k, n_rays, s, 21)
    grads = _zero_grads(shapes)
    rc = lib.dmnerf_mlp_bwd_bf16_noalloc(_p(packed), _p(acts), _p(d_raw), n_rays * s, k, _ptr_array(grads), _st())
    assert rc == 2, rc                                               # DMNERF_EUNSUPPORTED
    msg = _lib.last_error()
    assert "2048" in msg and "exceeds" in msg, msg
    torch.cuda.synchronize()
    assert all(float(g.abs().max()) == 0.0 for g in grads)           # refused before anything was launched
    rc = lib.dmnerf_mlp_bwd_bf16_noalloc(_p(None), _p(acts), _p(d_raw), 64, k, _ptr_array(grads), _st())
    assert rc == 1 and "null" in _lib.last_error().lower()           # DMNERF_EINVAL
