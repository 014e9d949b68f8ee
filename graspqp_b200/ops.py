"""torch.autograd bridges over the C-ABI kernels, for callers that keep the reference's autograd-driven loop
(`new_energy.sum().backward()`, scripts/fit.py:440).  Forward AND backward of every op are libgraspqp_b200.so kernels;
torch only owns the tensors and the tape.  CUDA tensors only — there is no CPU path."""
import ctypes

import torch

from . import _lib


def _st(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _chk(rc):
    _lib.check(rc)


def _f32c(t):
    if not t.is_cuda:
        raise RuntimeError("graspqp_b200 ops need CUDA tensors (no CPU fallback)")
    return t.detach().to(torch.float32).contiguous()


class SdfFn(torch.autograd.Function):
    """torchsdf._UnbatchedTriangleDistanceCuda (thirdparty/TorchSDF/torchsdf/sdf.py:42-68)."""

    @staticmethod
    def forward(ctx, points, face_vertices):
        p, fv = _f32c(points), _f32c(face_vertices)
        P, F = p.shape[0], fv.shape[0]
        d2 = torch.empty(P, device=p.device)
        sg = torch.empty(P, dtype=torch.int32, device=p.device)
        nr, cl = torch.empty(P, 3, device=p.device), torch.empty(P, 3, device=p.device)
        _chk(_lib.lib().gqp_sdf_forward(p.data_ptr(), fv.data_ptr(), P, F, d2.data_ptr(), sg.data_ptr(), nr.data_ptr(),
                                        cl.data_ptr(), _st(p)))
        ctx.save_for_backward(p, cl)
        ctx.mark_non_differentiable(sg, nr, cl)
        return d2, sg, nr, cl

    @staticmethod
    def backward(ctx, g, *_):
        p, cl = ctx.saved_tensors
        g = _f32c(g)
        out = torch.empty_like(p)
        _chk(_lib.lib().gqp_sdf_backward(g.data_ptr(), p.data_ptr(), cl.data_ptr(), p.shape[0], out.data_ptr(), _st(p)))
        return out, None


class FkFn(torch.autograd.Function):
    """hand_pose (Bt,9+dof) -> world link transforms (Bt,L,12)."""

    @staticmethod
    def forward(ctx, pose, hand):
        p = _f32c(pose)
        tf = torch.empty(p.shape[0], hand.n_links, 12, device=p.device)
        _chk(_lib.lib().gqp_fk_forward(hand.handle, p.data_ptr(), p.shape[0], tf.data_ptr(), _st(p)))
        ctx.hand = hand
        ctx.save_for_backward(p)
        return tf

    @staticmethod
    def backward(ctx, g):
        (p,) = ctx.saved_tensors
        g = _f32c(g)
        out = torch.empty_like(p)
        _chk(_lib.lib().gqp_fk_backward(ctx.hand.handle, p.data_ptr(), g.data_ptr(), None, p.shape[0], out.data_ptr(),
                                        _st(p)))
        return out, None


class ContactsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, link_tf, idx, hand):
        tf = _f32c(link_tf)
        ix = idx.detach().to(torch.int32).contiguous()
        B, n_c = ix.shape
        cp, cn = torch.empty(B, n_c, 3, device=tf.device), torch.empty(B, n_c, 3, device=tf.device)
        _chk(_lib.lib().gqp_contacts_forward(hand.handle, tf.data_ptr(), ix.data_ptr(), B, n_c, cp.data_ptr(),
                                             cn.data_ptr(), _st(tf)))
        ctx.hand, ctx.shape = hand, tf.shape
        ctx.save_for_backward(ix)
        return cp, cn

    @staticmethod
    def backward(ctx, gp, gn):
        (ix,) = ctx.saved_tensors
        B, n_c = ix.shape
        gp, gn = _f32c(gp), _f32c(gn)
        g = torch.zeros(ctx.shape, device=ix.device)
        _chk(_lib.lib().gqp_contacts_backward(ctx.hand.handle, ix.data_ptr(), gp.data_ptr(), gn.data_ptr(), B, n_c,
                                              g.data_ptr(), _st(g)))
        return g, None, None


class HandDistanceFn(torch.autograd.Function):
    """HandModel.cal_distance: x (Bt,P,3) per candidate -> signed depth (Bt,P), positive inside."""

    @staticmethod
    def forward(ctx, link_tf, x, hand):
        tf, xx = _f32c(link_tf), _f32c(x)
        B, P = xx.shape[0], xx.shape[1]
        lib = _lib.lib()
        dist = torch.empty(B, P, device=tf.device)
        arg = torch.empty(B, P, dtype=torch.int32, device=tf.device)
        dg = torch.empty(B, P, 3, device=tf.device)
        sc = torch.empty(int(lib.gqp_hand_distance_scratch_floats(B, 1, P, 1)), device=tf.device)
        _chk(lib.gqp_hand_distance_full(hand.handle, tf.data_ptr(), xx.data_ptr(), B, 1, P, 1, dist.data_ptr(),
                                        arg.data_ptr(), dg.data_ptr(), sc.data_ptr(), _st(tf)))
        ctx.hand = hand
        ctx.save_for_backward(tf, xx, arg, dg)
        return dist

    @staticmethod
    def backward(ctx, g):
        tf, xx, arg, dg = ctx.saved_tensors
        g = _f32c(g)
        out = torch.zeros_like(tf)
        B, P = xx.shape[0], xx.shape[1]
        _chk(_lib.lib().gqp_hand_distance_backward(ctx.hand.handle, tf.data_ptr(), xx.data_ptr(), B, 1, P, 1,
                                                   g.data_ptr(), arg.data_ptr(), dg.data_ptr(), out.data_ptr(), _st(tf)))
        return out, None, None


class SelfPenFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, link_tf, pose, hand):
        tf, p = _f32c(link_tf), _f32c(pose)
        B = p.shape[0]
        e = torch.empty(B, device=tf.device)
        g = torch.zeros_like(tf)
        _chk(_lib.lib().gqp_self_pen_joints(hand.handle, tf.data_ptr(), p.data_ptr(), B, e.data_ptr(), None, 1.0, 0.0,
                                            g.data_ptr(), None, _st(tf)))
        ctx.save_for_backward(g)
        return e

    @staticmethod
    def backward(ctx, ge):
        (g,) = ctx.saved_tensors
        return g * ge.to(g.dtype).view(-1, 1, 1), None, None


class ObjectDistanceFn(torch.autograd.Function):
    """ObjectModel.cal_distance: x (O*B, n, 3) -> distance (O*B, n) [differentiable], outward normals [not]."""

    @staticmethod
    def forward(ctx, x, obj):
        xx = _f32c(x)
        Bt, n = xx.shape[0], xx.shape[1]
        dist, sg = torch.empty(Bt, n, device=xx.device), torch.empty(Bt, n, dtype=torch.int32, device=xx.device)
        nr, cl = torch.empty(Bt, n, 3, device=xx.device), torch.empty(Bt, n, 3, device=xx.device)
        _chk(_lib.lib().gqp_object_distance_forward(obj.objset, obj.batch_size_each, n, xx.data_ptr(), obj.object_scale,
                                                    dist.data_ptr(), nr.data_ptr(), cl.data_ptr(), sg.data_ptr(), _st(xx)))
        ctx.scale = obj.object_scale
        ctx.save_for_backward(xx, cl, sg)
        ctx.mark_non_differentiable(nr)
        return dist, nr

    @staticmethod
    def backward(ctx, g, _gn):
        xx, cl, sg = ctx.saved_tensors
        g = _f32c(g)
        out = torch.empty_like(xx)
        _chk(_lib.lib().gqp_object_distance_backward(g.data_ptr(), xx.data_ptr(), cl.data_ptr(), sg.data_ptr(),
                                                     xx.shape[0] * xx.shape[1], ctx.scale, out.data_ptr(), _st(xx)))
        return out, None


class ForceClosureFn(torch.autograd.Function):
    """energy_fnc of MetricType.GRASPQP (registry.py:31-89): -> E_fc (Bt), per-contact force sums (Bt,n_c)."""

    @staticmethod
    def forward(ctx, contact_pts, contact_normals, cog, params):
        cp, cn, cg = _f32c(contact_pts), _f32c(contact_normals), _f32c(cog)
        B, n_c = cp.shape[0], cp.shape[1]
        lib = _lib.lib()
        e, lam = torch.empty(B, device=cp.device), torch.empty(B, n_c, device=cp.device)
        g = torch.empty_like(cp)
        zeros = torch.zeros(B, n_c, device=cp.device)
        ws = torch.empty((int(lib.gqp_fc_workspace_bytes(B, n_c, params.max_iter)) + 7) // 8, dtype=torch.float64,
                         device=cp.device)
        # cog is per candidate here (energy.py:35-42 passes (Bt,3)): n_obj = B "objects" of one candidate each
        _chk(lib.gqp_fc_solve(ctypes.byref(params), cp.data_ptr(), cn.data_ptr(), cg.data_ptr(), B, 1, n_c,
                              ws.data_ptr(), _st(cp)))
        _chk(lib.gqp_fc_finish(ctypes.byref(params), cp.data_ptr(), cn.data_ptr(), zeros.data_ptr(), cn.data_ptr(), None,
                               cg.data_ptr(), B, 1, n_c, ws.data_ptr(), 1, e.data_ptr(), None, lam.data_ptr(),
                               g.data_ptr(), None, None, _st(cp)))
        ctx.save_for_backward(g)
        ctx.mark_non_differentiable(lam)
        return e, lam

    @staticmethod
    def backward(ctx, ge, _gl):
        (g,) = ctx.saved_tensors
        return g * ge.to(g.dtype).view(-1, 1, 1), None, None, None


class SpanEnergyFn(torch.autograd.Function):
    """energy_fnc of MetricType.DEXGRASP / TDG (metrics/ops/dexgrasp.py, tdg.py) over gqp_span_energy -> E_fc (Bt)."""

    @staticmethod
    def forward(ctx, contact_pts, contact_normals, cog, params, dirs):
        cp, cn, cg = _f32c(contact_pts), _f32c(contact_normals), _f32c(cog)
        B, n_c = cp.shape[0], cp.shape[1]
        e, ed = torch.empty(B, device=cp.device), torch.empty(B, device=cp.device)
        g = torch.empty_like(cp)
        zeros = torch.zeros(B, n_c, device=cp.device)
        # cog is per candidate here (energy.py:35-42 passes (Bt,3)): n_obj = B "objects" of one candidate each
        _chk(_lib.lib().gqp_span_energy(ctypes.byref(params), cp.data_ptr(), cn.data_ptr(), zeros.data_ptr(), cn.data_ptr(),
                                        None, cg.data_ptr(), dirs.data_ptr() if dirs is not None else None,
                                        dirs.shape[0] if dirs is not None else 0, B, 1, n_c, e.data_ptr(), ed.data_ptr(),
                                        g.data_ptr(), None, _st(cp)))
        ctx.save_for_backward(g)
        return e

    @staticmethod
    def backward(ctx, ge):
        (g,) = ctx.saved_tensors
        return g * ge.to(g.dtype).view(-1, 1, 1), None, None, None, None
