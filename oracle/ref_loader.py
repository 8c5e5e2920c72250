"""Import the UNMODIFIED reference for checking and for the timed CPU baseline (test infrastructure, not product).

``load()`` puts the reference root at the front of ``sys.path`` -- ``oracle/_ref`` (staged by ``oracle/make_ref.sh``;
travels to the GPU box) or, in the authoring container, ``/root/reference`` -- after registering empty stand-ins for
the IO-only third-party modules its files import at module top and that this image lacks (imageio, lpips, skimage,
open3d, matplotlib, trimesh, h5py, configargparse: networks/tester.py:1-15, networks/manipulator.py:1-15,
tools/visualizer.py, config.py).  None of them is touched by the compute functions.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.
"""
from __future__ import annotations

import os
import sys
import types
from types import SimpleNamespace
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
IO_STUBS = ("imageio", "lpips", "skimage", "skimage.metrics", "skimage.measure", "open3d", "matplotlib",
            "matplotlib.pyplot", "trimesh", "h5py", "configargparse")


def reference_root() -> Optional[str]:
    cands = [os.environ.get("DMNERF_REFERENCE"), os.path.join(HERE, "_ref"), "/root/reference"]
    for c in cands:
        if c and os.path.isfile(os.path.join(c, "networks", "render.py")):
            return c
    return None


def stub_io_modules() -> None:
    for name in IO_STUBS:
        if name in sys.modules:
            continue
        try:
            __import__(name)
        except Exception:
            mod = types.ModuleType(name)
            mod.__dmnerf_stub__ = True
            sys.modules[name] = mod
    for parent, child in (("skimage", "metrics"), ("skimage", "measure"), ("matplotlib", "pyplot")):
        if not hasattr(sys.modules[parent], child):
            setattr(sys.modules[parent], child, sys.modules[f"{parent}.{child}"])


def functional_io_stubs(captured: Optional[dict] = None) -> dict:
    """Give the stand-ins just enough behaviour for ``render_test`` / ``manipulator_demo`` to run end to end with the
    metrics and the image writers neutralised (tester.py:43,96-98,139-149; manipulator.py:475-486): LPIPS -> 0, PSNR /
    SSIM -> 0, ``imageio.imwrite`` / ``cv2.imwrite`` record their arrays in ``captured[path]`` instead of writing."""
    import torch
    stub_io_modules()
    captured = {} if captured is None else captured

    class _LPIPS:
        def __init__(self, *a, **k):
            pass

        def to(self, *_):
            return self

        def __call__(self, a, b):
            return torch.zeros(1)

    def _imwrite(path, arr, *a, **k):
        captured[str(path)] = arr
        return True

    lp, sk, io = sys.modules["lpips"], sys.modules["skimage.metrics"], sys.modules["imageio"]
    if getattr(lp, "__dmnerf_stub__", False):
        lp.LPIPS = _LPIPS
    if getattr(sk, "__dmnerf_stub__", False):
        sk.peak_signal_noise_ratio = lambda *a, **k: 0.0
        sk.structural_similarity = lambda *a, **k: 0.0
    if getattr(io, "__dmnerf_stub__", False):
        io.imwrite = _imwrite
    import cv2
    cv2.imwrite = _imwrite
    return captured


def load(with_drivers: bool = False, keep_anomaly: bool = False) -> SimpleNamespace:
    """Returns the reference's hot-path modules (``model, render, helpers, pen, mani, evaluator``; with
    ``with_drivers`` also ``tester``).  The reference switches autograd anomaly detection on at import
    (networks/dm_nerf.py:5); it is switched off again unless ``keep_anomaly``."""
    import torch
    root = reference_root()
    if root is None:
        raise RuntimeError("reference not available: run oracle/make_ref.sh where /root/reference exists")
    stub_io_modules()
    if root not in sys.path:
        sys.path.insert(0, root)
    import networks.dm_nerf as r_model
    import networks.render as r_render
    import networks.helpers as r_helpers
    import networks.penalizer as r_pen
    import networks.evaluator as r_eval
    import networks.manipulator as r_mani
    ns = SimpleNamespace(root=root, model=r_model, render=r_render, helpers=r_helpers, pen=r_pen, mani=r_mani,
                         evaluator=r_eval)
    if with_drivers:
        import networks.tester as r_tester
        ns.tester = r_tester
    if not keep_anomaly:
        torch.autograd.set_detect_anomaly(False)
    for m in (r_model, r_render, r_helpers, r_pen, r_mani):
        assert os.path.abspath(m.__file__).startswith(os.path.abspath(root)), (m.__file__, root)
    return ns
