"""graspqp_b200 — B200-native GraspQP hot path: per-iteration grasp-energy evaluation + MALA* update.

Fast path:      graspqp_b200.engine.FitEngine      (whole iteration as ~20 kernel launches, no autograd, no host sync)
Reference API:  HandModel, ObjectModel, GraspSpanMetricFactory, calculate_energy, MalaStar, torchsdf.compute_sdf
                (same names / signatures as leggedrobotics/graspqp, autograd bridges over the same kernels)
Boundary:       include/graspqp_b200.h (C ABI of libgraspqp_b200.so).  No CPU fallback anywhere in this package.
"""
SDF_BACKEND = "B200"


def __getattr__(name):  # lazy: importing the package must not need torch/CUDA (build(), tools)
    import importlib
    table = {"HandModel": "hand_model", "ObjectModel": "object_model", "GraspSpanMetricFactory": "metrics",
             "calculate_energy": "energy", "MalaStar": "optimizer", "FitEngine": "engine"}
    if name in table:
        return getattr(importlib.import_module("." + table[name], __name__), name)
    raise AttributeError(name)
