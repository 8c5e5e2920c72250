"""dm_nerf_b200 -- B200-native (sm_100a) implementation of DM-NeRF's per-ray hot path.

``dm_nerf_b200.networks`` mirrors the reference's ``networks`` package call-for-call
(render.dm_nerf / render_train, dm_nerf.DM_NeRF / Embedder / get_embedder, helpers.sample_pdf /
z_val_sample / get_rays_k, penalizer.ins_penalizer, manipulator.manipulator / exchanger ...);
call ``dm_nerf_b200.install()`` before the reference's drivers import ``networks.*`` and they run on this path
(dropin.py: the reference's own package is patched name by name; its tester / metrics / IO stay untouched).
All computation is hand-written CUDA in ``libdmnerf_b200.so`` behind the C-ABI declared in
``include/dmnerf_b200.h``; there is no CPU / PyTorch fallback.
"""
from . import _lib  # noqa: F401
from .ops import set_mode, get_mode  # noqa: F401
from .dropin import install, uninstall  # noqa: F401

__version__ = "0.2.0"
