"""CPU-side checks of the boundary: the C-ABI library builds for sm_100a, loads, and exports every
symbol include/dmnerf_b200.h declares; the host mirror keeps the reference's names and state_dict."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dmnerf_b200.h")).read()
    return sorted(set(re.findall(r"\b(dmnerf_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_all_symbols():
    from dm_nerf_b200 import _lib
    _lib.build()
    lib = _lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in the header but not exported"
    assert set(syms) == set(_lib.EXPORTS), set(syms) ^ set(_lib.EXPORTS)
    assert lib.dmnerf_version() == 100 and lib.dmnerf_built_arch() == 100
    assert lib.dmnerf_packed_bytes(32, 0) > 698789 * 4
    assert lib.dmnerf_packed_bytes(500, 0) == -1


def test_state_dict_matches_reference_layout():
    import dmnerf_oracle as orc
    from dm_nerf_b200.networks.dm_nerf import DM_NeRF, get_embedder
    for k in (16, 32, 93):
        m = DM_NeRF(8, 256, 63, 27, [4], k)
        shapes = {n: tuple(v.shape) for n, v in m.state_dict().items()}
        assert shapes == orc.param_shapes(k)
        m.load_state_dict(orc.init_params(k, 1))
    pe, d = get_embedder(10, 0)
    assert d == 63 and get_embedder(4, 0)[1] == 27
    with pytest.raises(RuntimeError):
        DM_NeRF(4, 128, 63, 27, [2], 8)


def test_no_cpu_fallback():
    from dm_nerf_b200.networks.render import render_train
    with pytest.raises(RuntimeError, match="CUDA"):
        render_train(torch.zeros(2, 4, 9), torch.zeros(2, 4), torch.zeros(2, 3))


def test_surface_names():
    import dm_nerf_b200.networks as nw
    for mod, names in {"render": ["dm_nerf", "render_train", "render_rays", "raw2outputs"],
                       "helpers": ["sample_pdf", "z_val_sample", "get_rays_k", "get_select_full"],
                       "dm_nerf": ["DM_NeRF", "Embedder", "get_embedder"],
                       "penalizer": ["ins_penalizer", "emptiness_penalizer"],
                       "manipulator": ["manipulator", "manipulator_nerf", "manipulator_render", "exchanger"]}.items():
        for n in names:
            assert hasattr(getattr(nw, mod), n), (mod, n)
