"""graspqp_b200 — B200-native GraspQP hot path: per-iteration grasp-energy evaluation + MALA* update.

Fast path:      graspqp_b200.engine.FitEngine      (whole iteration as ~20 kernel launches, no autograd, no host sync)
Reference API:  HandModel, ObjectModel, GraspSpanMetricFactory, calculate_energy, MalaStar, AnnealingDexGraspNet,
                torchsdf.compute_sdf
                (same names / signatures as leggedrobotics/graspqp, autograd bridges over the same kernels)
Boundary:       include/graspqp_b200.h (C ABI of libgraspqp_b200.so).  No CPU fallback anywhere in this package.
"""
import os as _os

# The reference selects its SDF backend from the environment at import (hand_model.py:25-34, object_model.py:15-24):
# SDF_BACKEND in {TORCHSDF (default), WARP, KAOLIN}.  This package implements the TORCHSDF contract (same winner, sign
# heuristic and gradient, bit for bit) with its own kernels; "B200" is accepted as a synonym.  WARP / KAOLIN have
# different sign and tie semantics and are not built: selecting them fails loudly instead of silently changing results.
SDF_BACKEND = _os.environ.get("SDF_BACKEND", "TORCHSDF").upper()
if SDF_BACKEND not in ("TORCHSDF", "B200"):
    raise RuntimeError(f"SDF_BACKEND={SDF_BACKEND!r}: graspqp_b200 implements the TORCHSDF backend's contract only "
                       "(set SDF_BACKEND=TORCHSDF or unset it)")


def __getattr__(name):  # lazy: importing the package must not need torch/CUDA (build(), tools)
    import importlib
    table = {"HandModel": "hand_model", "ObjectModel": "object_model", "GraspSpanMetricFactory": "metrics",
             "calculate_energy": "energy", "MalaStar": "optimizer", "AnnealingDexGraspNet": "optimizer",
             "FitEngine": "engine", "get_hand_model": "hand_model", "AVAILABLE_HANDS": "hand_model"}
    if name in table:
        return getattr(importlib.import_module("." + table[name], __name__), name)
    raise AttributeError(name)
