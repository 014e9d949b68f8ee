"""BASELINE / TEST INFRASTRUCTURE — loads the reference's OWN Python modules for the hot path, unmodified, from a directory
holding the `graspqp` package and one holding `pytorch_kinematics` (in the build container: /root/reference; on the GPU
box: baseline/_ref/, put there by baseline/build_ref_gpu.sh).

The package __init__ files are bypassed (they import trimesh / pytorch3d / lxml, absent in this image) by pre-registering
empty package modules whose __path__ points at the directories.  Third-party modules that are NOT vendored in the
reference and not installed are replaced by stubs built from oracle/:

    qpth.qp.QPFunction        -> oracle.qp.QPFunction  (restatement of qpth 0.0.18's batched PDIPM; PARITY UNPINNED)
    roma.special_gramschmidt  -> oracle.hand.special_gramschmidt (restatement of roma 1.5.4;        PARITY UNPINNED)
"""
import importlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_loaded = {}


def _pkg(name, path):
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


def load(graspqp_dir, pk_dir):
    """-> dict of reference modules (optimizer, energy, span, registry, qp_solver, transforms, pk_chain, pk_frame, pk_tf)"""
    if _loaded:
        return _loaded
    from oracle import hand as ohand
    from oracle import qp as oqp

    # ---- stubs of the un-vendored third-party modules ----
    class QPFunction:  # qpth.qp.QPFunction(verbose, maxIter, eps)(Q, p, G, h, A, b)
        def __init__(self, eps=1e-12, verbose=0, notImprovedLim=3, maxIter=20, **kw):
            self.eps, self.maxIter, self.notImprovedLim = eps, maxIter, notImprovedLim

        def __call__(self, Q, p, G, h, A, b):
            assert A.numel() == 0 and b.numel() == 0, "the hot path has no equality constraints"
            assert (self.eps, self.maxIter) == (5e-2, 12), "oracle.qp is specialised to the reference's settings"
            if G.ndim == 3:
                G = G[0]
            return oqp.QPFunction.apply(Q, p, G.to(Q.dtype), h)

    qpth = types.ModuleType("qpth")
    qpth.qp = types.ModuleType("qpth.qp")
    qpth.qp.QPFunction = QPFunction
    sys.modules.setdefault("qpth", qpth)
    sys.modules.setdefault("qpth.qp", qpth.qp)
    roma = types.ModuleType("roma")
    roma.special_gramschmidt = ohand.special_gramschmidt
    sys.modules.setdefault("roma", roma)

    # ---- the reference's own packages, without their __init__ ----
    g = graspqp_dir
    for name, path in [("graspqp", g), ("graspqp.core", g + "/core"), ("graspqp.metrics", g + "/metrics"),
                       ("graspqp.metrics.ops", g + "/metrics/ops"), ("graspqp.metrics.solver", g + "/metrics/solver"),
                       ("graspqp.utils", g + "/utils")]:
        _pkg(name, path)
    _pkg("pytorch_kinematics", pk_dir)
    imp = importlib.import_module
    _loaded.update(
        optimizer=imp("graspqp.core.optimizer"), energy=imp("graspqp.core.energy"),
        span=imp("graspqp.metrics.ops.span"), registry=imp("graspqp.metrics.ops.registry"),
        qp_solver=imp("graspqp.metrics.solver.qp_solver"), transforms=imp("graspqp.utils.transforms"),
        pk_tf=imp("pytorch_kinematics.transforms"), pk_frame=imp("pytorch_kinematics.frame"),
        pk_chain=imp("pytorch_kinematics.chain"))
    return _loaded
