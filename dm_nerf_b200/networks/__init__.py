"""Shadow of the reference's ``networks`` package for the per-ray hot path."""
from . import dm_nerf, helpers, render, penalizer, manipulator, evaluator  # noqa: F401
