"""Force-closure metric registry — graspqp/src/graspqp/metrics/ops/registry.py:92-140 for MetricType.GRASPQP.
The callable has the reference contract (core/energy.py:35-42):
    energy_fnc(contact_pts, contact_normals, sdf, cog, with_solution=True, svd_gain=0.1) -> (E_fc (Bt), x (Bt,n_c) | None)
Other metric types (dexgrasp, tdg, handle) are other energy_types: out of this path's scope."""
import enum

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


class GraspSpanMetricFactory:
    class MetricType(enum.Enum):
        GRASPQP = "graspqp"

    @staticmethod
    def create(metric_type, solver_kwargs=None, **kwargs):
        if metric_type != GraspSpanMetricFactory.MetricType.GRASPQP:
            raise NotImplementedError(f"{metric_type}: only energy_type=graspqp is on the accelerated path")
        return SpanMetricWrapper(**(solver_kwargs or {}))
