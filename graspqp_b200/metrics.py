"""Force-closure metric registry — graspqp/src/graspqp/metrics/ops/registry.py:92-140 for MetricType.GRASPQP.
The callable has the reference contract (core/energy.py:35-42):
    energy_fnc(contact_pts, contact_normals, sdf, cog, with_solution=True, svd_gain=0.1) -> (E_fc (Bt), x (Bt,n_c) | None)
MetricType.DEXGRASP and TDG are the reference's other energy_types (scripts/fit.py:344-347) over
gqp_span_energy; the scipy solvers (GRASPQP_SCIPY, GRASPQP_EUCLIDIAN_SCIPY) are CPU code outside this path."""
import enum

import numpy as np
import torch

from . import _lib, ops


class SpanMetricWrapper:
    """registry.py:31-89 + span.py:313-415 + qp_solver.py:60-134, one kernel launch family (csrc/gqp_fc.cu)."""

    def __init__(self, friction=0.2, max_limit=20.0, n_cone_vecs=4, torque_weight=5.0, values_gain=2.0):
        if n_cone_vecs != 4:
            raise NotImplementedError("the sm_100a kernel implements the reference default n_cone_vecs=4")
        self.friction, self.max_limit, self.torque_weight, self.values_gain = friction, max_limit, torque_weight, values_gain

    def __call__(self, contact_pts, contact_normals, sdf=None, cog=None, with_solution=False, svd_gain=0.1, **kwargs):
        p = _lib.GqpFcParams(friction=self.friction, torque_weight=self.torque_weight, lower=1.0,
                             upper=self.max_limit + 1.0, ridge=1e-4, svd_gain=svd_gain, values_gain=self.values_gain,
                             values_offset=1e-2, eps=5e-2, max_iter=12, w_fc=1.0, w_dis=0.0)
        e, lam = ops.ForceClosureFn.apply(contact_pts, contact_normals, cog, p)
        return (e, lam) if with_solution else (e, None)

    forward = __call__


class DexgraspSpanMetric:
    """metrics/ops/dexgrasp.py:39-69: E = |sum n_i|^2 + tw^2 |sum n_i x (p_i - cog)|^2; returns (E, ones) with_solution."""

    def __call__(self, contact_pts, contact_normals, cog=None, torque_weight=0.0, with_solution=False, **kwargs):
        p = _lib.GqpSpanParams(kind=_lib.SPAN_DEXGRASP, torque_weight=float(torque_weight), friction=0.2, obb_length=0.2,
                               scale=100.0, enable_density=1, w_fc=1.0, w_dis=0.0)
        e = ops.SpanEnergyFn.apply(contact_pts, contact_normals, cog, p, None)
        return (e, torch.ones_like(contact_pts[..., 0])) if with_solution else e

    forward = __call__


class TDGSpanMetric:
    """metrics/ops/tdg.py:217-239: 100 * TDGEnergy(miu 0.2, box length 0.2, density weights) over 1000 unit directions drawn
    with np.random.randn at construction (tdg.py:44-48, 148-150: from numpy's global state, as the reference)."""

    def __init__(self, device="cuda", directions=None):
        if directions is None:
            d = np.random.randn(1000, 3)
            directions = d / (np.linalg.norm(d, axis=-1)[:, None] + 1e-8)
        self.directions = torch.as_tensor(np.ascontiguousarray(directions, np.float32)).to(device).contiguous()

    def __call__(self, contact_pts, contact_normals, cog=None, torque_weight=0.0, with_solution=False, **kwargs):
        p = _lib.GqpSpanParams(kind=_lib.SPAN_TDG, torque_weight=0.0, friction=0.2, obb_length=0.2, scale=100.0,
                               enable_density=1, w_fc=1.0, w_dis=0.0)
        return ops.SpanEnergyFn.apply(contact_pts, contact_normals, cog, p, self.directions.to(contact_pts.device)), None

    forward = __call__


class GraspSpanMetricFactory:
    class MetricType(enum.Enum):  # registry.py:94-99
        DEXGRASP = 1
        TDG = 2
        GRASPQP = 3
        GRASPQP_SCIPY = 4
        GRASPQP_EUCLIDIAN_SCIPY = 5

    @staticmethod
    def create(metric_type, solver_kwargs=None, **kwargs):
        T = GraspSpanMetricFactory.MetricType
        if metric_type == T.DEXGRASP:
            return DexgraspSpanMetric()
        if metric_type == T.TDG:
            return TDGSpanMetric(**kwargs)
        if metric_type == T.GRASPQP:
            return SpanMetricWrapper(**(solver_kwargs or {}))
        raise NotImplementedError(f"{metric_type}: the scipy solvers are CPU code outside the accelerated path")
