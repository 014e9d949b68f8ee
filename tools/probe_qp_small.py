"""Diagnostic (test infrastructure): small-QP regime (n_c = 4, two parallel pads => nearly rank-deficient grasp matrix).
Three evaluations of E_fc on identical inputs:
  K  the kernel (csrc/gqp_fc.cu through the host emulator: 6x6 Woodbury form, fp64)
  O  oracle/qp.py in torch.float64 (full KKT through LU of the 2n x 2n Schur complement)
  M  the same algorithm as O in 40-digit arithmetic (mpmath) — the rounding-free iterate
Prints |K-M|/|M| and |O-M|/|M| per case.  Usage: python tools/probe_qp_small.py [n_cases] [seed]"""
import ctypes
import os
import sys

import mpmath as mp
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from graspqp_b200 import _lib, build  # noqa: E402
from oracle import qp as oqp  # noqa: E402

mp.mp.dps = 40


def pads_inputs(B, n_c, seed, jitter=1e-3):
    """contacts on two parallel pads (normals ~ +-x), as the Robotiq 2F-140 produces at n_c = 4"""
    rng = np.random.default_rng(seed)
    side = np.where(np.arange(n_c) % 2 == 0, 1.0, -1.0)
    nrm = np.zeros((B, n_c, 3))
    nrm[..., 0] = side
    nrm += rng.normal(size=nrm.shape) * jitter
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    cp = np.zeros((B, n_c, 3))
    cp[..., 0] = -side * 0.03
    cp[..., 1:] = rng.normal(size=(B, n_c, 2)) * 0.02
    cog = rng.normal(size=(B, 3)) * 0.005
    return cp.astype(np.float32), nrm.astype(np.float32), cog.astype(np.float32)


def mp_pdipm(Fm, lo=1.0, hi=21.0, ridge=1e-4, eps=5e-2, max_iter=12, stop_at=None):
    """oracle.qp.pdipm_forward for ONE problem in mpmath; Newton systems solved in the reduced (n x n) form, which is
    algebraically the same step.  Returns the per-iteration residuals and the list of iterates (x, s, z)."""
    n = Fm.cols
    Q = Fm.T * Fm + mp.eye(n) * mp.mpf(ridge)
    hi, lo = mp.mpf(hi), mp.mpf(lo)

    def solve(du, dl, rx, rsu, rsl, rzu, rzl):
        A = Q.copy()
        r = mp.matrix(n, 1)
        for j in range(n):
            A[j, j] += du[j] + dl[j]
            r[j] = -rx[j] + (rsu[j] - du[j] * rzu[j]) - (rsl[j] - dl[j] * rzl[j])
        dx = mp.lu_solve(A, r)
        dzu = [-rsu[j] + du[j] * rzu[j] + du[j] * dx[j] for j in range(n)]
        dzl = [-rsl[j] + dl[j] * rzl[j] - dl[j] * dx[j] for j in range(n)]
        dsu = [-rzu[j] - dx[j] for j in range(n)]
        dsl = [-rzl[j] + dx[j] for j in range(n)]
        return [dx[j] for j in range(n)], dsu, dsl, dzu, dzl

    one, zero = [mp.mpf(1)] * n, [mp.mpf(0)] * n
    x, su, sl, zu, zl = solve(one, one, zero, zero, zero, [-hi] * n, [lo] * n)
    ms, mz = min(su + sl), min(zu + zl)
    if ms < 0:
        su, sl = [v - (ms - 1) for v in su], [v - (ms - 1) for v in sl]
    if mz < 0:
        zu, zl = [v - (mz - 1) for v in zu], [v - (mz - 1) for v in zl]
    m = 2 * n
    out = []

    def step(v, dv):
        c = [-a / b for a, b in zip(v, dv) if b < 0]
        return min(c) if c else mp.mpf(10) ** 300

    for it in range(max_iter):
        Qx = Q * mp.matrix(x)
        rx = [zu[j] - zl[j] + Qx[j] for j in range(n)]
        rzu = [x[j] + su[j] - hi for j in range(n)]
        rzl = [-x[j] + sl[j] + lo for j in range(n)]
        sz = sum(su[j] * zu[j] + sl[j] * zl[j] for j in range(n))
        mu = abs(sz) / m
        res = mp.sqrt(sum(v * v for v in rzu + rzl)) + mp.sqrt(sum(v * v for v in rx)) + m * mu
        out.append((res, list(x), list(su), list(sl), list(zu), list(zl)))
        if it == max_iter - 1:
            break
        du = [zu[j] / su[j] for j in range(n)]
        dl = [zl[j] / sl[j] for j in range(n)]
        dx, dsu, dsl, dzu, dzl = solve(du, dl, rx, zu, zl, rzu, rzl)
        a1 = min(step(zu + zl, dzu + dzl), step(su + sl, dsu + dsl), mp.mpf(1))
        t3 = sum((su[j] + a1 * dsu[j]) * (zu[j] + a1 * dzu[j]) + (sl[j] + a1 * dsl[j]) * (zl[j] + a1 * dzl[j])
                 for j in range(n))
        sig = (t3 / sz) ** 3
        rs2u = [(-mu * sig + dsu[j] * dzu[j]) / su[j] for j in range(n)]
        rs2l = [(-mu * sig + dsl[j] * dzl[j]) / sl[j] for j in range(n)]
        cx, csu, csl, czu, czl = solve(du, dl, zero, rs2u, rs2l, zero, zero)
        dx = [a + b for a, b in zip(dx, cx)]
        dsu = [a + b for a, b in zip(dsu, csu)]
        dsl = [a + b for a, b in zip(dsl, csl)]
        dzu = [a + b for a, b in zip(dzu, czu)]
        dzl = [a + b for a, b in zip(dzl, czl)]
        a2 = min(mp.mpf("0.999") * min(step(zu + zl, dzu + dzl), step(su + sl, dsu + dsl)), mp.mpf(1))
        x = [x[j] + a2 * dx[j] for j in range(n)]
        su = [su[j] + a2 * dsu[j] for j in range(n)]
        sl = [sl[j] + a2 * dsl[j] for j in range(n)]
        zu = [zu[j] + a2 * dzu[j] for j in range(n)]
        zl = [zl[j] + a2 * dzl[j] for j in range(n)]
    return out


def mp_energy(F64, stop, svd_gain=0.1):
    """E_fc of one problem from the best iterate among iterations 0..stop (qpth keeps the best-residual iterate)"""
    Fm = mp.matrix(F64.tolist())
    its = mp_pdipm(Fm)
    k = min(stop, len(its) - 1)
    best = min(range(k + 1), key=lambda i: (its[i][0], i))
    x = mp.matrix(its[best][1])
    Fx = Fm * x
    value = sum(v * v for v in Fx) / 2
    det = mp.det(Fm * Fm.T)
    svd = mp.sqrt(det) ** (mp.mpf(1) / 6) if det > 0 else mp.mpf(0)
    return float(2 * (value + mp.mpf("0.01")) * mp.exp(-mp.mpf(svd_gain) * svd)), best, [float(i[0]) for i in its]


def run_kernel(lib, cp, nrm, cog, n_c):
    B = len(cp)
    P = _lib.GqpFcParams(friction=0.2, torque_weight=5.0, lower=1.0, upper=21.0, ridge=1e-4, svd_gain=0.1,
                         values_gain=2.0, values_offset=1e-2, eps=5e-2, max_iter=12, w_fc=1.0, w_dis=0.0)
    ws = np.zeros((int(lib.gqp_fc_workspace_bytes(B, n_c, 12)) + 7) // 8, np.float64)
    e, ed, lam = np.zeros(B, np.float32), np.zeros(B, np.float32), np.zeros((B, n_c), np.float32)
    zeros = np.zeros((B, n_c), np.float32)
    stop = np.zeros(1, np.int32)
    p = lambda a: a.ctypes.data
    assert lib.gqp_fc_solve(ctypes.byref(P), p(cp), p(nrm), p(cog), B, 1, n_c, p(ws), None) == 0
    assert lib.gqp_fc_finish(ctypes.byref(P), p(cp), p(nrm), p(zeros), p(nrm), None, p(cog), B, 1, n_c, p(ws), 1, p(e),
                             p(ed), p(lam), None, None, p(stop), None) == 0, lib.gqp_last_error()
    return e, int(stop[0])


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    n_c = 4
    lib = _lib.bind(ctypes.CDLL(build.build_emu()), strict=False)
    cp, nrm, cog = pads_inputs(B, n_c, seed)
    eK, stop = run_kernel(lib, cp, nrm, cog, n_c)
    t = lambda a: torch.tensor(a, dtype=torch.float64)
    tr = []
    F = oqp.grasp_matrix(t(cp), t(nrm), t(cog))
    eO, _, info = oqp.graspqp_energy(t(cp), t(nrm), t(cog))
    print("kernel exit iteration", stop, " cond(F F') per case:",
          ["%.1e" % c for c in torch.linalg.cond(F @ F.mT).tolist()])
    for b in range(B):
        eM, best, res = mp_energy(F[b].numpy(), stop)
        print(f"case {b}: M={eM:.9g}  |K-M|/M={abs(eK[b] - eM) / eM:.2e}  |O-M|/M={abs(eO[b].item() - eM) / eM:.2e}"
              f"  best it {best}")
