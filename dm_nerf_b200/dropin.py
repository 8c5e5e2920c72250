"""Drop-in binding: make the reference's own drivers run on the B200 path.

The reference has no plugin layer; its boundary is the Python call surface of ``networks/*``.  ``install()`` imports
the reference's OWN ``networks`` package (it must be importable: run the drivers from the reference checkout, or pass
``reference_root``) and replaces, name by name, exactly the hot-path callables of SURVEY 8(a):

    networks.dm_nerf      Embedder, get_embedder, DM_NeRF                       (dm_nerf.py:8-106)
    networks.render       dm_nerf, render_train                                 (render.py:6-96)
    networks.helpers      sample_pdf, z_val_sample, get_rays_k, get_rays,
                          get_select_full, get_select_crop                      (helpers.py:24-155)
    networks.penalizer    emptiness_penalizer, ins_penalizer                    (penalizer.py:5-62)
    networks.manipulator  exchanger, manipulator_render, manipulator_nerf,
                          manipulator                                           (manipulator.py:18-205)
    networks.evaluator    ins_criterion                                         (evaluator.py:19-74)

Everything else stays the reference's: ``networks.tester.render_test``, ``networks.evaluator.{to8b, ins_eval,
img2mse, mse2psnr, ...}``, ``networks.manipulator.{manipulator_eval, manipulator_demo}``, ``config``, loaders, tools.
Modules that bound a replaced name with ``from ... import`` (``networks.tester``, ``networks.manipulator``,
``tools.mesh_generator``, ``config``, the ``train_*`` / ``test_*`` scripts) are re-bound: at once when already
imported, otherwise right after their import through a meta-path hook.

When the reference is not importable at all, ``install()`` falls back to aliasing ``dm_nerf_b200.networks`` as
``networks`` (hot-path modules only) so that code written against the call surface still binds.
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.util
import os
import sys
from typing import Dict, Optional

_HOT = {
    "dm_nerf": ("Embedder", "get_embedder", "DM_NeRF"),
    "render": ("dm_nerf", "render_train"),
    "helpers": ("sample_pdf", "z_val_sample", "get_rays_k", "get_rays", "get_select_full", "get_select_crop"),
    "penalizer": ("emptiness_penalizer", "ins_penalizer"),
    "manipulator": ("exchanger", "manipulator_render", "manipulator_nerf", "manipulator"),
    "evaluator": ("ins_criterion",),
}
# imported eagerly by install(): torch + numpy (+ scipy) only.  networks.manipulator imports image-IO packages at
# module top (manipulator.py:3-15), so it is patched when (if) the driver imports it.
_EAGER = ("dm_nerf", "render", "helpers", "penalizer", "evaluator")

_replaced: Dict[int, object] = {}      # id(reference object) -> replacement
_originals = []                        # keeps the replaced objects alive (ids stay unique)
_state = {"mode": None, "hook": None}


def _ours(name: str):
    return importlib.import_module(f"dm_nerf_b200.networks.{name}")


def _patch_hot_module(short: str, mod) -> None:
    ours = _ours(short)
    for attr in _HOT[short]:
        new = getattr(ours, attr)
        old = getattr(mod, attr, None)
        if old is new:
            continue
        if old is not None:
            _replaced[id(old)] = new
            _originals.append(old)
        setattr(mod, attr, new)
    mod.__dmnerf_b200__ = True


def _rebind(mod) -> None:
    """Swap every module-level binding that still points at a replaced reference object."""
    d = getattr(mod, "__dict__", None)
    if not d:
        return
    for k, v in list(d.items()):
        new = _replaced.get(id(v))
        if new is not None:
            d[k] = new


def _after_import(mod) -> None:
    name = getattr(mod, "__name__", "")
    if name.startswith("networks."):
        short = name.split(".", 1)[1]
        if short in _HOT:
            _patch_hot_module(short, mod)
    _rebind(mod)


class _PostImportHook(importlib.abc.MetaPathFinder):
    """Runs ``_after_import`` on reference modules imported after ``install()``."""

    def __init__(self):
        self._busy = set()

    @staticmethod
    def _watched(fullname: str) -> bool:
        top = fullname.split(".", 1)[0]
        return top in ("networks", "tools", "datasets", "config") or top.startswith(("train_", "test_"))

    def find_spec(self, fullname, path=None, target=None):
        if not self._watched(fullname) or fullname in self._busy:
            return None
        self._busy.add(fullname)
        try:
            spec = importlib.util.find_spec(fullname)
        except (ImportError, ValueError):
            spec = None
        finally:
            self._busy.discard(fullname)
        if spec is None or spec.loader is None or not hasattr(spec.loader, "exec_module"):
            return None
        inner = spec.loader

        class _Loader(importlib.abc.Loader):
            def create_module(self, s):
                return inner.create_module(s) if hasattr(inner, "create_module") else None

            def exec_module(self, module):
                inner.exec_module(module)
                _after_import(module)

            def __getattr__(self, item):            # get_source / get_filename ... for tracebacks and inspect
                return getattr(inner, item)

        spec.loader = _Loader()
        return spec


def _reference_importable(reference_root: Optional[str]) -> bool:
    if reference_root is not None:
        if not os.path.isfile(os.path.join(reference_root, "networks", "render.py")):
            raise RuntimeError(f"install: {reference_root} does not hold the reference's networks/ package")
        if reference_root not in sys.path:
            sys.path.insert(0, reference_root)
    mod = sys.modules.get("networks")
    if mod is not None:
        ours = sys.modules.get("dm_nerf_b200.networks")
        return mod is not ours
    try:
        spec = importlib.util.find_spec("networks.render")
    except (ImportError, ValueError):
        return False
    return spec is not None and "dm_nerf_b200" not in (spec.origin or "")


def install(reference_root: Optional[str] = None, keep_anomaly_detection: bool = False) -> str:
    """Bind the reference's drivers to the B200 path; returns ``"patched"`` (reference package patched in place) or
    ``"aliased"`` (reference absent: ``networks`` is this package's mirror).

    The reference switches autograd anomaly detection on when ``networks.dm_nerf`` is imported (dm_nerf.py:5); that
    is a debugging aid which records a Python traceback per autograd node, so it is switched off again unless
    ``keep_anomaly_detection``."""
    if _reference_importable(reference_root):
        import torch
        for short in _EAGER:
            _patch_hot_module(short, importlib.import_module(f"networks.{short}"))
        if _state["hook"] is None:
            _state["hook"] = _PostImportHook()
            sys.meta_path.insert(0, _state["hook"])
        for name, mod in list(sys.modules.items()):
            if mod is not None and _PostImportHook._watched(name):
                _after_import(mod)
        main = sys.modules.get("__main__")
        if main is not None:
            _rebind(main)
        if not keep_anomaly_detection:
            torch.autograd.set_detect_anomaly(False)
        _state["mode"] = "patched"
    else:
        pkg = importlib.import_module("dm_nerf_b200.networks")
        sys.modules["networks"] = pkg
        for short in ("dm_nerf", "render", "helpers", "penalizer", "manipulator", "evaluator"):
            sys.modules[f"networks.{short}"] = _ours(short)
        _state["mode"] = "aliased"
    return _state["mode"]


def uninstall() -> None:
    """Undo ``install()`` (tests): restore the reference's objects, drop the hook / the aliases."""
    if _state["mode"] == "patched":
        back = {id(new): old for old in _originals for new in [_replaced.get(id(old))] if new is not None}
        for name, mod in list(sys.modules.items()):
            if mod is None or not _PostImportHook._watched(name):
                continue
            for k, v in list(getattr(mod, "__dict__", {}).items()):
                if id(v) in back:
                    mod.__dict__[k] = back[id(v)]
        if _state["hook"] in sys.meta_path:
            sys.meta_path.remove(_state["hook"])
        _state["hook"] = None
        _replaced.clear()
        _originals.clear()
    elif _state["mode"] == "aliased":
        for name in [n for n in sys.modules if n == "networks" or n.startswith("networks.")]:
            if "dm_nerf_b200" in getattr(sys.modules[name], "__name__", ""):
                del sys.modules[name]
    _state["mode"] = None
