"""dm_nerf_b200 -- B200-native (sm_100a) implementation of DM-NeRF's per-ray hot path.

``dm_nerf_b200.networks`` mirrors the reference's ``networks`` package call-for-call
(render.dm_nerf / render_train, dm_nerf.DM_NeRF / Embedder / get_embedder, helpers.sample_pdf /
z_val_sample / get_rays_k, penalizer.ins_penalizer, manipulator.manipulator / exchanger ...);
put ``dm_nerf_b200.install()`` before the reference's imports and its drivers run on this path.
All computation is hand-written CUDA in ``libdmnerf_b200.so`` behind the C-ABI declared in
``include/dmnerf_b200.h``; there is no CPU / PyTorch fallback.
"""
from . import _lib  # noqa: F401
from .ops import set_mode, get_mode  # noqa: F401

__version__ = "0.1.0"


def install() -> None:
    """Alias this package's ``networks`` modules as top-level ``networks.*`` so that the reference's
    drivers (``from networks.render import dm_nerf`` ...) bind to the B200 path."""
    import importlib
    import sys
    pkg = importlib.import_module("dm_nerf_b200.networks")
    sys.modules["networks"] = pkg
    for name in ("dm_nerf", "render", "helpers", "penalizer", "manipulator"):
        sys.modules[f"networks.{name}"] = importlib.import_module(f"dm_nerf_b200.networks.{name}")
