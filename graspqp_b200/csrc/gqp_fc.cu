// Q1-Q4 + E_dis — the force-closure metric of energy_type=graspqp and the contact-distance term, forward and analytic
// backward in ONE launch, a warp per candidate.
//   friction cone  : metrics/ops/span.py:263-295            grasp matrix F (6 x n): span.py:341-346
//   QP             : metrics/solver/qp_solver.py:60-134 -> qpth.qp.QPFunction(maxIter=12, eps=5e-2) (qpth 0.0.18)
//   E_fc           : metrics/ops/registry.py:82-87, svd term span.py:402        E_dis: core/energy.py:25-28
//
// The QP is  min 1/2 x'(F'F + ridge I)x  s.t.  lower <= x <= upper  (n = 4 n_c <= 64 variables, 2n inequalities).
// qpth runs a primal-dual interior point method whose Newton systems it solves through a 2n x 2n Schur complement in
// fp32 LU.  The same Newton steps are computed here in reduced form:  (F'F + M) dx = r  with M = ridge + d_u + d_l
// diagonal, i.e. by the Woodbury identity  dx = M^-1 r - M^-1 F'(I6 + F M^-1 F')^-1 F M^-1 r : one 6x6 SPD system per
// solve.  Every lane owns <= 2 variables (its F columns, x, slacks, duals: registers); the 6x6 matrix is reduced across
// the warp through shared memory and factorised redundantly by every lane.  All QP arithmetic is fp64 (B200 runs fp64
// at half the fp32 rate; the stage is ~25 % of an iteration at config 2), so the iterates are the exact-arithmetic iterates of
// qpth's algorithm rather than one fp32 realisation of them.
// Implicit differentiation (qpth QPFunction.backward): one more solve with d = clamp(lam)/clamp(slack).
#include "gqp_common.cuh"
#include "graspqp_b200.h"

// One warp per block, 16 blocks per SM (128 registers, 136 B of spills).  Measured on B200, fc_solve at Bt = 8192 /
// 32768: 4 warps x 3 blocks (168 registers) 313 / 1174 us; 1 x 10-12 (168) 250 / 931; 1 x 14-16 (128) 234 / 848;
// 1 x 20 (96) 253 / 928; 1 x 24 (80) 311 / 1116.
#ifndef FC_WARPS
#define FC_WARPS 1
#endif
#ifndef FC_MIN_BLOCKS
#define FC_MIN_BLOCKS 16
#endif
#define FC_RED_K 29  // widest reduction (odd: conflict-free column sums)
#define FC_SM_PER_WARP (32 * FC_RED_K + 64 + 64 * 3)

struct FcParamsDev {
    double mu_f, c0, tw, lo, hi, ridge, svd_gain, values_gain, values_offset;
    int max_iter;
    double w_fc, w_dis;
    float eps;
};

// all-reduce K doubles across the warp: v[k] <- op over lanes.  Narrow reductions (K <= 6) are shuffle butterflies;
// wide ones go through shared memory, each of K lanes summing one column in 4 independent chains.
// G = 32: the warp holds one candidate; G = 16: two candidates, one per half-warp (n <= 32 variables), every reduction
// stays inside its half.
template <int K, bool IS_MIN, int G>
__device__ __forceinline__ void warp_allreduce(double *sm, double (&v)[K], int lane) {
    if (K <= 6) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) {
                double t = __shfl_xor_sync(GQP_FULL, v[k], o);
                v[k] = IS_MIN ? (t < v[k] ? t : v[k]) : v[k] + t;
            }
        }
        return;
    }
    const int S = K | 1, gl = lane & (G - 1), base = lane - gl;
#pragma unroll
    for (int k = 0; k < K; ++k) sm[lane * S + k] = v[k];
    __syncwarp();
    for (int col = gl; col < K; col += G) {
        const double *c0 = sm + base * S + col;
        double a0 = c0[0], a1 = c0[S], a2 = c0[2 * S], a3 = c0[3 * S];
#pragma unroll
        for (int r = 4; r < G; r += 4) {
            double t0 = c0[r * S], t1 = c0[(r + 1) * S], t2 = c0[(r + 2) * S], t3 = c0[(r + 3) * S];
            if (IS_MIN) {
                a0 = t0 < a0 ? t0 : a0; a1 = t1 < a1 ? t1 : a1; a2 = t2 < a2 ? t2 : a2; a3 = t3 < a3 ? t3 : a3;
            } else {
                a0 += t0; a1 += t1; a2 += t2; a3 += t3;
            }
        }
        double a = IS_MIN ? fmin(fmin(a0, a1), fmin(a2, a3)) : (a0 + a1) + (a2 + a3);
        sm[32 * FC_RED_K + base * 2 + col] = a;
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < K; ++k) v[k] = sm[32 * FC_RED_K + base * 2 + k];
    __syncwarp();
}

// Reciprocal and reciprocal square root of a positive normal double: the hardware seed (about 20 bits) and two Newton
// steps, without the library routines' special-case branches (a dozen of them sit on the critical path of every
// interior-point iteration).  Within an ulp or two of the correctly rounded result; NaN for negative input.
__device__ __forceinline__ double fc_rcp(double x) {
#ifdef GQP_EMU
    return 1.0 / x;
#else
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
#endif
}
__device__ __forceinline__ double fc_rsqrt(double x) {
#ifdef GQP_EMU
    return rsqrt(x);
#else
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-(x * y), y, 1.0);
    y = fma(0.5 * y, e, y);
    e = fma(-(x * y), y, 1.0);
    return fma(0.5 * y, e, y);
#endif
}

// packed lower-triangular index
__device__ __forceinline__ int tri_ix(int i, int j) { return i * (i + 1) / 2 + j; }

// in-place Cholesky of a packed 6x6 SPD matrix (lower), A = L L'; diagonal entries are stored INVERTED.
// Every loop has the constant trip count 6 with the triangular condition as a predicate: the compiler then unrolls all
// of them and the matrix stays in registers (with data-dependent bounds it kept one loop rolled and the 21 entries went
// to local memory, indexed dynamically).
template <bool FAST>
__device__ __forceinline__ void chol6(double *A) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double s = A[tri_ix(j, j)];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k < j) s -= A[tri_ix(j, k)] * A[tri_ix(j, k)];
        double inv = FAST ? fc_rsqrt(s) : rsqrt(s);  // the diagonal holds 1/L_jj: the solves multiply instead of dividing
        A[tri_ix(j, j)] = inv;
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i > j) {
                double t = A[tri_ix(i, j)];
#pragma unroll
                for (int k = 0; k < 6; ++k)
                    if (k < j) t -= A[tri_ix(i, k)] * A[tri_ix(j, k)];
                A[tri_ix(i, j)] = t * inv;
            }
    }
}
__device__ __forceinline__ void chol6_solve(const double *L, double *y) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double t = y[i];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k < i) t -= L[tri_ix(i, k)] * y[k];
        y[i] = t * L[tri_ix(i, i)];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double t = y[i];
#pragma unroll
        for (int k = 0; k < 6; ++k)
            if (k > i) t -= L[tri_ix(k, i)] * y[k];
        y[i] = t * L[tri_ix(i, i)];
    }
}

struct Dir {
    double dx[2], dsu[2], dsl[2], dzu[2], dzl[2];
};

// step length to the boundary (qpth get_step): min over entries with dv < 0 of -v/dv, +inf otherwise.  The running
// minimum is kept as a fraction num/den (den > 0) and compared by cross-multiplication: one division per lane.
struct Frac {
    double num, den;
};
__device__ __forceinline__ void step_min(Frac &f, double v, double dv) {
    if (dv < 0.0 && v * f.den < f.num * (-dv)) {
        f.num = v;
        f.den = -dv;
    }
}
__device__ __forceinline__ double step_lane(const double *zu, const double *zl, const double *su, const double *sl,
                                            const Dir &d, const bool *valid) {
    Frac f = {1e300, 1.0};
#pragma unroll
    for (int v = 0; v < 2; ++v)
        if (valid[v]) {
            step_min(f, zu[v], d.dzu[v]); step_min(f, zl[v], d.dzl[v]);
            step_min(f, su[v], d.dsu[v]); step_min(f, sl[v], d.dsl[v]);
        }
    return f.num * fc_rcp(f.den);
}

// Workspace of the split form (gqp_fc_solve -> gqp_fc_finish).  The interior-point pass keeps, per candidate, every
// iterate that improved its best residual (x and the backward's d = clamp(z)/clamp(s) per bound: 3n doubles) and which
// iterate was best as of each iteration; batch-wide it keeps, per iteration, the maximum best residual and whether
// any candidate improved — exactly what qpth's exit test reads.  The finish pass derives the exit iteration from those
// 2 x max_iter words and completes every candidate from its stored iterate: one IPM pass instead of two.
struct FcWorkspace {
    unsigned *gmax;     // (FC_MAX_ITER) float bits of max_b best_res_b after iteration it
    unsigned *gimp;     // (FC_MAX_ITER) != 0 if any candidate improved at iteration it
    uint8_t *best_idx;  // (Bt, FC_MAX_ITER) index of the candidate's best iterate as of iteration it
    double *snap;       // (Bt, max_iter, 3, n)
};
#define FC_MAX_ITER 64
#define FC_WS_HEADER 512
__host__ __device__ static inline size_t fc_align256(size_t v) { return (v + 255) / 256 * 256; }
static inline FcWorkspace fc_workspace_view(void *ws, int Bt) {
    FcWorkspace w;
    unsigned char *b = (unsigned char *)ws;
    w.gmax = (unsigned *)b;
    w.gimp = w.gmax + FC_MAX_ITER;
    w.best_idx = b + FC_WS_HEADER;
    w.snap = (double *)(b + FC_WS_HEADER + fc_align256((size_t)Bt * FC_MAX_ITER));
    return w;
}

// qpth's exit test (pdipm/batch.py forward): leave at the first iteration where no candidate improved for 3 consecutive
// iterations or the batch maximum of the best residuals is below eps
__device__ __forceinline__ int fc_batch_stop(const unsigned *gmax, const unsigned *gimp, int max_iter, float eps) {
    int not_improved = 0;
    for (int it = 0; it < max_iter; ++it) {
        const float m2 = __int_as_float((int)gmax[it]);
        if (it > 0) not_improved = gimp[it] ? 0 : not_improved + 1;
        if (not_improved == 3 || m2 < eps) return it;
    }
    return max_iter;
}

enum { FC_FUSED = 0, FC_SOLVE = 1, FC_FINISH = 2 };

template <int MODE, int G>
__global__ void __launch_bounds__(FC_WARPS * 32, FC_MIN_BLOCKS)
fc_dis_kernel(FcParamsDev P, const float *__restrict__ cpts, const float *__restrict__ cnrm_hand,
              const float *__restrict__ obj_dist, const float *__restrict__ obj_nrm,
              const float *__restrict__ obj_dgrad, const float *__restrict__ cog, int Bt, int batch_each, int n_c,
              float *__restrict__ e_fc, float *__restrict__ e_dis, float *__restrict__ lam_out,
              float *__restrict__ grad_cpts, float *__restrict__ grad_cnrm, const int32_t *__restrict__ stop_iter,
              float *__restrict__ trace, FcWorkspace ws, int batch_stop, int32_t *__restrict__ stop_out) {
    __shared__ double smem[FC_WARPS][FC_SM_PER_WARP + 64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gl = lane & (G - 1), grp = lane / G, NG = 32 / G;  // lane within its candidate's group, group, groups/warp
    double *sm = smem[warp];
    double *sm4 = sm + 32 * FC_RED_K + 64;  // (64,4) per-variable scratch
    const int n = 4 * n_c, m = 2 * n;
    int stop = P.max_iter;
    if (MODE == FC_FUSED && stop_iter) stop = *stop_iter;
    if (MODE == FC_FINISH) {
        if (batch_stop) stop = fc_batch_stop(ws.gmax, ws.gimp, P.max_iter, P.eps);
        if (stop_out && blockIdx.x == 0 && threadIdx.x == 0) *stop_out = stop;
    }
    const double INF = 1e300;

    for (int wb = blockIdx.x * FC_WARPS + warp; wb * NG < Bt; wb += gridDim.x * FC_WARPS) {
        const bool live = wb * NG + grp < Bt;  // an odd batch leaves the last warp's second half idle: it computes the
        const int b = live ? wb * NG + grp : Bt - 1;  // last candidate again and stores nothing
        const int o = b / batch_each;
        // ---- Q1/Q2: friction-cone edges and grasp-matrix columns of this lane's variables ----
        double F[2][6];
        bool valid[2];
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            const int j = gl + G * v;
            valid[v] = j < n;
#pragma unroll
            for (int i = 0; i < 6; ++i) F[v][i] = 0.0;
            if (valid[v]) {
                const int c = j >> 2, k = j & 3;
                const size_t q = ((size_t)b * n_c + c) * 3;
                double nx = obj_nrm[q], ny = obj_nrm[q + 1], nz = obj_nrm[q + 2];
                double b1x = 0.57735026918962576, b1y = b1x, b1z = b1x;
                double dp = (b1x * nx + b1y * ny + b1z * nz) * 1.0 / (sqrt(nx * nx + ny * ny + nz * nz) + 1e-6);
                if (dp > 0.9) b1y -= 2.0;
                double t1x = ny * b1z - nz * b1y, t1y = nz * b1x - nx * b1z, t1z = nx * b1y - ny * b1x;
                double tx = t1x, ty = t1y, tz = t1z;
                if (k & 1) {  // t2 = n x t1
                    tx = ny * t1z - nz * t1y; ty = nz * t1x - nx * t1z; tz = nx * t1y - ny * t1x;
                }
                double sg = (k & 2) ? -P.mu_f : P.mu_f;
                double fx = (sg * tx + P.c0 * nx) * 0.25, fy = (sg * ty + P.c0 * ny) * 0.25,
                       fz = (sg * tz + P.c0 * nz) * 0.25;
                double rx = (double)cpts[q] - cog[3 * o], ry = (double)cpts[q + 1] - cog[3 * o + 1],
                       rz = (double)cpts[q + 2] - cog[3 * o + 2];
                F[v][0] = fx; F[v][1] = fy; F[v][2] = fz;
                F[v][3] = (ry * fz - rz * fy) * P.tw;
                F[v][4] = (rz * fx - rx * fz) * P.tw;
                F[v][5] = (rx * fy - ry * fx) * P.tw;
            }
        }
        // ---- Q3: qpth's PDIPM, Newton systems in Woodbury form ----
        double Minv[2], L6[21];
        auto factor = [&](const double *du, const double *dl) {
            double a[21];
#pragma unroll
            for (int i = 0; i < 21; ++i) a[i] = 0.0;
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                Minv[v] = valid[v] ? fc_rcp(P.ridge + du[v] + dl[v]) : 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i)
#pragma unroll
                    for (int j2 = 0; j2 <= i; ++j2) a[tri_ix(i, j2)] += Minv[v] * F[v][i] * F[v][j2];
            }
            warp_allreduce<21, false, G>(sm, a, lane);
#pragma unroll
            for (int i = 0; i < 6; ++i) a[tri_ix(i, i)] += 1.0;
            chol6<true>(a);  // I + F M^-1 F': pivots >= 1
#pragma unroll
            for (int i = 0; i < 21; ++i) L6[i] = a[i];
        };
        // also returns y = (I + F M^-1 F')^-1 F M^-1 r, which equals F dx
        auto solve = [&](const double *du, const double *dl, const double *rx, const double *rsu, const double *rsl,
                         const double *rzu, const double *rzl, Dir &d, double (&t)[6]) {
            double r[2];
#pragma unroll
            for (int i = 0; i < 6; ++i) t[i] = 0.0;
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                r[v] = -rx[v] + (rsu[v] - du[v] * rzu[v]) - (rsl[v] - dl[v] * rzl[v]);
#pragma unroll
                for (int i = 0; i < 6; ++i) t[i] += F[v][i] * r[v] * Minv[v];
            }
            warp_allreduce<6, false, G>(sm, t, lane);
            chol6_solve(L6, t);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                double fty = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) fty += F[v][i] * t[i];
                double dx = Minv[v] * (r[v] - fty);
                d.dx[v] = dx;
                d.dzu[v] = -rsu[v] + du[v] * rzu[v] + du[v] * dx;
                d.dzl[v] = -rsl[v] + dl[v] * rzl[v] - dl[v] * dx;
                d.dsu[v] = -rzu[v] - dx;
                d.dsl[v] = -rzl[v] + dx;
            }
        };
        double x[2], su[2], sl[2], zu[2], zl[2];
        double fxv[6];  // F x, carried along: F dx comes out of every Newton solve, no reduction needed
        double bx[2] = {0, 0}, bsu[2] = {1, 1}, bsl[2] = {1, 1}, bzu[2] = {1, 1}, bzl[2] = {1, 1};
        const double zero2[2] = {0.0, 0.0};
        if (MODE != FC_FINISH) {
        {   // starting point: KKT solve with d = 1, then shift slacks / duals to >= 1 (qpth pdipm forward)
            const double one2[2] = {1.0, 1.0};
            double rzu[2] = {-P.hi, -P.hi}, rzl[2] = {P.lo, P.lo};
            factor(one2, one2);
            Dir d;
            solve(one2, one2, zero2, zero2, zero2, rzu, rzl, d, fxv);  // x = dx, so F x = y
            double mn[2] = {INF, INF};
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                x[v] = d.dx[v]; su[v] = d.dsu[v]; sl[v] = d.dsl[v]; zu[v] = d.dzu[v]; zl[v] = d.dzl[v];
                if (valid[v]) {
                    mn[0] = fmin(mn[0], fmin(su[v], sl[v]));
                    mn[1] = fmin(mn[1], fmin(zu[v], zl[v]));
                }
            }
            warp_allreduce<2, true, G>(sm, mn, lane);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                if (mn[0] < 0.0) { su[v] -= mn[0] - 1.0; sl[v] -= mn[0] - 1.0; }
                if (mn[1] < 0.0) { zu[v] -= mn[1] - 1.0; zl[v] -= mn[1] - 1.0; }
            }
        }
        double best_res = INF;
        int best_it = 0;
        const double inv_m = 1.0 / m;
        for (int it = 0; it < P.max_iter; ++it) {
            double rx[2], rzu[2], rzl[2], du[2], dl[2], isu[2], isl[2], acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                double ftfx = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) ftfx += F[v][i] * fxv[i];
                rx[v] = zu[v] - zl[v] + ftfx + P.ridge * x[v];
                rzu[v] = x[v] + su[v] - P.hi;
                rzl[v] = -x[v] + sl[v] + P.lo;
                isu[v] = fc_rcp(su[v]);
                isl[v] = fc_rcp(sl[v]);
                du[v] = zu[v] * isu[v];
                dl[v] = zl[v] * isl[v];
                if (valid[v]) {
                    acc[0] += su[v] * zu[v] + sl[v] * zl[v];
                    acc[1] += rzu[v] * rzu[v] + rzl[v] * rzl[v];
                    acc[2] += rx[v] * rx[v];
                } else {
                    rx[v] = rzu[v] = rzl[v] = 0.0;
                    du[v] = dl[v] = 1.0;
                }
            }
            warp_allreduce<3, false, G>(sm, acc, lane);
            const double mu = fabs(acc[0]) * inv_m;
            const double res = acc[1] * fc_rsqrt(fmax(acc[1], 1e-300)) + acc[2] * fc_rsqrt(fmax(acc[2], 1e-300)) + m * mu;
            if (trace && gl == 0 && live) trace[(size_t)b * P.max_iter + it] = (float)res;
            const bool improved = it == 0 || res < best_res;
            if (improved) {
                best_res = res;
                best_it = it;
                if (MODE == FC_SOLVE && live) {
                    double *sn = ws.snap + ((size_t)b * P.max_iter + it) * 3 * n;
#pragma unroll
                    for (int v = 0; v < 2; ++v)
                        if (valid[v]) {
                            const int j = gl + G * v;
                            sn[j] = x[v];
                            sn[n + j] = fmax(zu[v], 1e-8) * (su[v] >= 1e-8 ? isu[v] : 1e8);
                            sn[2 * n + j] = fmax(zl[v], 1e-8) * (sl[v] >= 1e-8 ? isl[v] : 1e8);
                        }
                } else if (MODE != FC_SOLVE) {
#pragma unroll
                    for (int v = 0; v < 2; ++v) { bx[v] = x[v]; bsu[v] = su[v]; bsl[v] = sl[v]; bzu[v] = zu[v]; bzl[v] = zl[v]; }
                }
            }
            if (MODE == FC_SOLVE && gl == 0 && live) {
                ws.best_idx[(size_t)b * FC_MAX_ITER + it] = (uint8_t)best_it;
                // batch statistics: the maximum only grows and the flag only sets, so a plain read filters most atomics
                const unsigned bits = (unsigned)__float_as_int((float)best_res);
                if (*(volatile unsigned *)(ws.gmax + it) < bits) atomicMax(ws.gmax + it, bits);
                if (improved && it > 0 && *(volatile unsigned *)(ws.gimp + it) == 0u) atomicOr(ws.gimp + it, 1u);
            }
            if (it >= stop || it == P.max_iter - 1) break;  // the step after the last examined iterate is never used
            factor(du, dl);
            Dir aff;
            double ya[6], yc[6];
            solve(du, dl, rx, zu, zl, rzu, rzl, aff, ya);  // rs = z
            double al[1] = {step_lane(zu, zl, su, sl, aff, valid)};
            warp_allreduce<1, true, G>(sm, al, lane);
            const double a1 = fmin(al[0], 1.0);
            double t3[1] = {0.0};
#pragma unroll
            for (int v = 0; v < 2; ++v)
                if (valid[v])
                    t3[0] += (su[v] + a1 * aff.dsu[v]) * (zu[v] + a1 * aff.dzu[v]) +
                             (sl[v] + a1 * aff.dsl[v]) * (zl[v] + a1 * aff.dzl[v]);
            warp_allreduce<1, false, G>(sm, t3, lane);
            const double ratio = t3[0] * fc_rcp(acc[0]);
            const double sig = ratio * ratio * ratio;
            double rs2u[2], rs2l[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                rs2u[v] = valid[v] ? (-mu * sig + aff.dsu[v] * aff.dzu[v]) * isu[v] : 0.0;
                rs2l[v] = valid[v] ? (-mu * sig + aff.dsl[v] * aff.dzl[v]) * isl[v] : 0.0;
            }
            Dir cor;
            solve(du, dl, zero2, rs2u, rs2l, zero2, zero2, cor, yc);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                aff.dx[v] += cor.dx[v]; aff.dsu[v] += cor.dsu[v]; aff.dsl[v] += cor.dsl[v];
                aff.dzu[v] += cor.dzu[v]; aff.dzl[v] += cor.dzl[v];
            }
            al[0] = step_lane(zu, zl, su, sl, aff, valid);
            warp_allreduce<1, true, G>(sm, al, lane);
            const double a2 = fmin(0.999 * al[0], 1.0);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                x[v] += a2 * aff.dx[v]; su[v] += a2 * aff.dsu[v]; sl[v] += a2 * aff.dsl[v];
                zu[v] += a2 * aff.dzu[v]; zl[v] += a2 * aff.dzl[v];
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) fxv[i] += a2 * (ya[i] + yc[i]);
        }
        }  // MODE != FC_FINISH
        if (MODE == FC_SOLVE) continue;
        if (!e_fc && !e_dis && !lam_out && !grad_cpts && !grad_cnrm) continue;  // trace-only pass
        double bdu[2] = {1.0, 1.0}, bdl[2] = {1.0, 1.0};
        if (MODE == FC_FINISH) {
            const int kk = stop < P.max_iter ? stop : P.max_iter - 1;
            const int bi = ws.best_idx[(size_t)b * FC_MAX_ITER + kk];
            const double *sn = ws.snap + ((size_t)b * P.max_iter + bi) * 3 * n;
#pragma unroll
            for (int v = 0; v < 2; ++v)
                if (valid[v]) {
                    const int j = gl + G * v;
                    bx[v] = sn[j]; bdu[v] = sn[n + j]; bdl[v] = sn[2 * n + j];
                }
        } else {
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                bdu[v] = fmax(bzu[v], 1e-8) / fmax(bsu[v], 1e-8);
                bdl[v] = fmax(bzl[v], 1e-8) / fmax(bsl[v], 1e-8);
            }
        }
        // ---- value, svd term, E_fc (qp_solver.py:126, span.py:402, registry.py:82-87) ----
        double red[27];
#pragma unroll
        for (int i = 0; i < 27; ++i) red[i] = 0.0;
#pragma unroll
        for (int v = 0; v < 2; ++v)
            if (valid[v]) {
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    red[i] += F[v][i] * bx[v];
#pragma unroll
                    for (int j2 = 0; j2 <= i; ++j2) red[6 + tri_ix(i, j2)] += F[v][i] * F[v][j2];
                }
            }
        warp_allreduce<27, false, G>(sm, red, lane);
        double Fx[6], Lg[21];
        double value = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) { Fx[i] = red[i]; value += 0.5 * red[i] * red[i]; }
#pragma unroll
        for (int i = 0; i < 21; ++i) Lg[i] = red[6 + i];
        double piv = 1.0;  // smallest pivot of the factorisation relative to its diagonal entry of F F'
        {
            double gd[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) gd[i] = Lg[tri_ix(i, i)];
            chol6<false>(Lg);  // F F' = Lg Lg'  ->  prod(svdvals(F)) = prod(diag Lg)
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const double li = Lg[tri_ix(i, i)];  // 1 / L_ii
                const double r = 1.0 / (li * li * gd[i]);
                piv = (r == r) ? fmin(piv, r) : 0.0;
            }
        }
        const bool need_qr = piv < 1e-6;
        if (__any_sync(GQP_FULL, need_qr)) {  // uniform over the warp: the other half (G = 16) runs along, result unused
            // Nearly rank-deficient grasp matrix (contacts on one line, duplicated contact indices: n_c = 4 on parallel
            // pads).  Cholesky of F F' squares the condition number and resolves sigma_min/sigma_max only down to 1e-8,
            // where svdvals (span.py:402) resolves 1e-16, and prod(sigma)^(1/6) magnifies exactly that.  Re-factorise
            // F' = Q R by classical Gram-Schmidt with re-orthogonalisation (CGS2) over the warp: R'R = F F' and the
            // entries of R carry an absolute error of eps * |F|, as svdvals' do.  Warp-uniform, rare branch.
            double a[2][6], R[6][6];
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 6; ++i) a[v][i] = valid[v] ? F[v][i] : 0.0;
#pragma unroll
            for (int j = 0; j < 6; ++j) {
#pragma unroll
                for (int k = 0; k < 6; ++k) R[k][j] = 0.0;
#pragma unroll
                for (int pass = 0; pass < 2; ++pass) {
                    double c[6];
#pragma unroll
                    for (int k = 0; k < 6; ++k) c[k] = (k < j) ? a[0][k] * a[0][j] + a[1][k] * a[1][j] : 0.0;
                    warp_allreduce<6, false, G>(sm, c, lane);
#pragma unroll
                    for (int k = 0; k < 6; ++k)
                        if (k < j) {
                            a[0][j] -= c[k] * a[0][k];
                            a[1][j] -= c[k] * a[1][k];
                            R[k][j] += c[k];
                        }
                }
                double nn[1] = {a[0][j] * a[0][j] + a[1][j] * a[1][j]};
                warp_allreduce<1, false, G>(sm, nn, lane);
                const double rjj = sqrt(nn[0]);
                R[j][j] = rjj;
                const double inv = rjj > 0.0 ? 1.0 / rjj : 0.0;
                a[0][j] *= inv;
                a[1][j] *= inv;
            }
            if (need_qr) {
#pragma unroll
                for (int i = 0; i < 6; ++i)
#pragma unroll
                    for (int k = 0; k <= i; ++k) Lg[tri_ix(i, k)] = (k == i) ? 1.0 / R[i][i] : R[k][i];
            }
        }
        double logdet = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) logdet -= log(Lg[tri_ix(i, i)]);  // diagonal is stored inverted
        const bool rank_ok = logdet == logdet && logdet > -1e300;  // NaN / -inf when F F' is numerically singular
        // Gradient of the svd term: d prod(sigma)^(1/6) / dF = svd/6 * (F F')^-1 F grows like 1/sigma_min.  When F is rank
        // deficient (sigma_min / sigma_max below fp32 resolution: all contacts on one line, duplicated contact indices at
        // n_c = 4) the reference's own fp32 value of it is rounding noise of size ~1e3..1e5 in an arbitrary direction, and
        // the rounding-free value is ~1e14+: it overflowed fp32 here, and ONE inf/NaN gradient row poisons (g^2).mean(0)
        // of the whole batch (optimizer.py:231-236 turns the NaN into ema = 0, i.e. a 1e6 x step for every candidate —
        // seen at config 4, 131072 Panda candidates, within 50 iterations).  The term's gradient is dropped there.
        double rmin = 1e300, rmax = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const double ri = 1.0 / Lg[tri_ix(i, i)];
            rmin = fmin(rmin, ri);
            rmax = fmax(rmax, ri);
        }
        const bool svd_grad_ok = rank_ok && rmin > 1e-7 * rmax;
        const double svd = rank_ok ? exp(logdet / 6.0) : 0.0;
        const double ex = exp(-P.svd_gain * svd);
        const double Efc = P.values_gain * (value + P.values_offset) * ex;
        // ---- E_dis (energy.py:25-28) on lanes < n_c ----
        double edis = 0.0, gpd[3] = {0.0, 0.0, 0.0};
        if (gl < n_c) {
            const size_t q = ((size_t)b * n_c + gl) * 3;
            double ox = obj_nrm[q], oy = obj_nrm[q + 1], oz = obj_nrm[q + 2];
            double a = 1.0 + (ox * cnrm_hand[q] + oy * cnrm_hand[q + 1] + oz * cnrm_hand[q + 2]);
            double dd = obj_dist[(size_t)b * n_c + gl];
            double ea = exp(a);
            edis = ea * fabs(dd);
            if (grad_cnrm && live) {
                double g = P.w_dis * edis;
                grad_cnrm[q] = (float)(g * ox); grad_cnrm[q + 1] = (float)(g * oy); grad_cnrm[q + 2] = (float)(g * oz);
            }
            double gd = P.w_dis * ea * (dd > 0.0 ? 1.0 : (dd < 0.0 ? -1.0 : 0.0));
            if (obj_dgrad) {
                gpd[0] = gd * obj_dgrad[q]; gpd[1] = gd * obj_dgrad[q + 1]; gpd[2] = gd * obj_dgrad[q + 2];
            }
        }
        double es[1] = {edis};
        warp_allreduce<1, false, G>(sm, es, lane);
        if (gl == 0 && live) {
            if (e_fc) e_fc[b] = (float)Efc;
            if (e_dis) e_dis[b] = (float)es[0];
        }
        // ---- backward: d(w_fc E_fc)/dF, implicit differentiation of the QP (qpth QPFunction.backward) ----
        double gF[2][6];
#pragma unroll
        for (int v = 0; v < 2; ++v)
#pragma unroll
            for (int i = 0; i < 6; ++i) gF[v][i] = 0.0;
        if (grad_cpts) {
            const double gv = P.w_fc * P.values_gain * ex;         // dE/dvalue
            const double gs = -P.w_fc * P.svd_gain * Efc;          // dE/dsvd
            double du[2], dl[2], rxb[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                double ftfx = 0.0;
#pragma unroll
                for (int i = 0; i < 6; ++i) ftfx += F[v][i] * Fx[i];
                rxb[v] = valid[v] ? gv * ftfx : 0.0;  // dl/dx
                du[v] = bdu[v];
                dl[v] = bdl[v];
            }
            factor(du, dl);
            Dir bd;
            double fdx[6];  // F dx
            solve(du, dl, rxb, zero2, zero2, zero2, zero2, bd, fdx);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                if (!valid[v]) continue;
                double pin[6];  // (F F')^-1 F_j
#pragma unroll
                for (int i = 0; i < 6; ++i) pin[i] = F[v][i];
                if (svd_grad_ok) chol6_solve(Lg, pin);
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    gF[v][i] = (gv * Fx[i] + fdx[i]) * bx[v] + Fx[i] * bd.dx[v] +
                               (svd_grad_ok ? gs * (svd / 6.0) * pin[i] : 0.0);
            }
        }
        // ---- per-contact sums: force solution (span.py:414) and d/d(contact point) through tau = tw (r x f) ----
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            const int j = gl + G * v + 32 * grp;  // G = 16: the second candidate's variables sit in slots 32..63
            if (gl + G * v < 64 / NG) {
                double fx = F[v][0], fy = F[v][1], fz = F[v][2], gx = gF[v][3], gy = gF[v][4], gz = gF[v][5];
                sm4[4 * j + 0] = valid[v] ? P.tw * (fy * gz - fz * gy) : 0.0;
                sm4[4 * j + 1] = valid[v] ? P.tw * (fz * gx - fx * gz) : 0.0;
                sm4[4 * j + 2] = valid[v] ? P.tw * (fx * gy - fy * gx) : 0.0;
                sm4[4 * j + 3] = valid[v] ? bx[v] : 0.0;
            }
        }
        __syncwarp();
        if (gl < n_c && live) {
            double s[4] = {0.0, 0.0, 0.0, 0.0};
            for (int k = 0; k < 4; ++k)
                for (int i = 0; i < 4; ++i) s[i] += sm4[4 * (4 * gl + k + 32 * grp) + i];
            const size_t q = ((size_t)b * n_c + gl) * 3;
            if (grad_cpts) {
                grad_cpts[q] = (float)(s[0] + gpd[0]);
                grad_cpts[q + 1] = (float)(s[1] + gpd[1]);
                grad_cpts[q + 2] = (float)(s[2] + gpd[2]);
            }
            if (lam_out) lam_out[(size_t)b * n_c + gl] = (float)s[3];
        }
        __syncwarp();
    }
}

// batch-wide exit test of qpth's loop, replayed from the residual trace (one block of 256 threads)
__global__ void fc_stop_iter_kernel(const float *__restrict__ trace, int Bt, int max_iter, float eps,
                                    int32_t *__restrict__ stop_iter) {
    __shared__ int s_improved;
    __shared__ float s_max[256];
    __shared__ int s_stop;
    int not_improved = 0;
    if (threadIdx.x == 0) s_stop = max_iter;
    for (int it = 0; it < max_iter; ++it) {
        if (threadIdx.x == 0) s_improved = 0;
        __syncthreads();
        float mx = -1e30f;
        int imp = 0;
        for (int b = threadIdx.x; b < Bt; b += blockDim.x) {
            float best = trace[(size_t)b * max_iter];
            for (int k = 1; k < it; ++k) best = fminf(best, trace[(size_t)b * max_iter + k]);
            float r = trace[(size_t)b * max_iter + it];
            if (it > 0 && r < best) imp = 1;
            mx = fmaxf(mx, it > 0 ? fminf(best, r) : r);
        }
        s_max[threadIdx.x] = mx;
        if (imp) atomicAdd(&s_improved, 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            float m2 = -1e30f;
            for (int i = 0; i < (int)blockDim.x; ++i) m2 = fmaxf(m2, s_max[i]);
            if (it > 0) not_improved = s_improved > 0 ? 0 : not_improved + 1;
            if ((not_improved == 3 || m2 < eps) && s_stop == max_iter) s_stop = it;
        }
        __syncthreads();
        if (s_stop != max_iter) break;
    }
    if (threadIdx.x == 0) *stop_iter = s_stop;
}

extern "C" int gqp_fc_dis_energy(const GqpFcParams *p, const float *cpts, const float *cnrm_hand,
                                 const float *obj_dist, const float *obj_nrm, const float *obj_dgrad, const float *cog,
                                 int n_obj, int batch_each, int n_c, float *e_fc, float *e_dis, float *lam,
                                 float *grad_cpts, float *grad_cnrm, const int32_t *qp_stop_iter, float *qp_trace,
                                 void *stream) {
    if (!p || n_obj < 0 || batch_each < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_dis_energy: bad argument");
    if (n_c < 2 || n_c > 16) return gqp_set_error(GQP_ERR_LIMIT, "gqp_fc_dis_energy: n_c must be in [2,16], got %d", n_c);
    if (p->max_iter < 1 || p->max_iter > 64) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_dis_energy: max_iter out of range");
    const int Bt = n_obj * batch_each;
    if (Bt == 0) return GQP_OK;
    if (!cpts || !cnrm_hand || !obj_dist || !obj_nrm || !cog)
        return gqp_set_error(GQP_ERR_ARG, "gqp_fc_dis_energy: null pointer");
    FcParamsDev P;
    P.mu_f = p->friction;
    P.c0 = sqrt(1.0 - (double)p->friction * (double)p->friction);
    P.tw = p->torque_weight; P.lo = p->lower; P.hi = p->upper; P.ridge = p->ridge; P.svd_gain = p->svd_gain;
    P.values_gain = p->values_gain; P.values_offset = p->values_offset; P.max_iter = p->max_iter;
    P.w_fc = p->w_fc; P.w_dis = p->w_dis; P.eps = p->eps;
    // n = 4 n_c <= 32 variables: two candidates per warp (half-warp reductions), else one
    if (n_c <= 8) {
        const int grid = ((Bt + 1) / 2 + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_FUSED, 16>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, cnrm_hand, obj_dist,
               obj_nrm, obj_dgrad, cog, Bt, batch_each, n_c, e_fc, e_dis, lam, grad_cpts, grad_cnrm, qp_stop_iter,
               qp_trace, FcWorkspace{}, 0, (int32_t *)nullptr);
    } else {
        const int grid = (Bt + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_FUSED, 32>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, cnrm_hand, obj_dist,
               obj_nrm, obj_dgrad, cog, Bt, batch_each, n_c, e_fc, e_dis, lam, grad_cpts, grad_cnrm, qp_stop_iter,
               qp_trace, FcWorkspace{}, 0, (int32_t *)nullptr);
    }
    GQP_CHECK_LAUNCH("gqp_fc_dis_energy");
    return GQP_OK;
}

extern "C" uint64_t gqp_fc_workspace_bytes(int Bt, int n_c, int max_iter) {
    if (Bt < 0 || n_c < 0 || max_iter < 0) return 0;
    return FC_WS_HEADER + fc_align256((size_t)Bt * FC_MAX_ITER) + (size_t)Bt * max_iter * 3 * 4 * n_c * sizeof(double) + 256;
}

static int fc_check(const char *who, const GqpFcParams *p, int n_obj, int batch_each, int n_c, FcParamsDev &P) {
    if (!p || n_obj < 0 || batch_each < 0) return gqp_set_error(GQP_ERR_ARG, "%s: bad argument", who);
    if (n_c < 2 || n_c > 16) return gqp_set_error(GQP_ERR_LIMIT, "%s: n_c must be in [2,16], got %d", who, n_c);
    if (p->max_iter < 1 || p->max_iter > FC_MAX_ITER) return gqp_set_error(GQP_ERR_ARG, "%s: max_iter out of range", who);
    P.mu_f = p->friction;
    P.c0 = sqrt(1.0 - (double)p->friction * (double)p->friction);
    P.tw = p->torque_weight; P.lo = p->lower; P.hi = p->upper; P.ridge = p->ridge; P.svd_gain = p->svd_gain;
    P.values_gain = p->values_gain; P.values_offset = p->values_offset; P.max_iter = p->max_iter;
    P.w_fc = p->w_fc; P.w_dis = p->w_dis; P.eps = p->eps;
    return GQP_OK;
}

extern "C" int gqp_fc_solve(const GqpFcParams *p, const float *cpts, const float *obj_nrm, const float *cog, int n_obj,
                            int batch_each, int n_c, void *workspace, void *stream) {
    FcParamsDev P;
    if (int rc = fc_check("gqp_fc_solve", p, n_obj, batch_each, n_c, P)) return rc;
    const int Bt = n_obj * batch_each;
    if (Bt == 0) return GQP_OK;
    if (!cpts || !obj_nrm || !cog || !workspace) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_solve: null pointer");
    if (((uintptr_t)workspace & 7) != 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_solve: workspace must be 8-byte aligned");
    FcWorkspace ws = fc_workspace_view(workspace, Bt);
#ifndef GQP_EMU
    if (cudaMemsetAsync(workspace, 0, FC_WS_HEADER, (cudaStream_t)stream) != cudaSuccess)
        return gqp_set_error(GQP_ERR_CUDA, "gqp_fc_solve: memset failed");
#else
    memset(workspace, 0, FC_WS_HEADER);
#endif
    // n = 4 n_c <= 32 variables: two candidates per warp (half-warp reductions), else one
    if (n_c <= 8) {
        const int grid = ((Bt + 1) / 2 + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_SOLVE, 16>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, (const float *)nullptr,
               (const float *)nullptr, obj_nrm, (const float *)nullptr, cog, Bt, batch_each, n_c, (float *)nullptr,
               (float *)nullptr, (float *)nullptr, (float *)nullptr, (float *)nullptr, (const int32_t *)nullptr,
               (float *)nullptr, ws, 0, (int32_t *)nullptr);
    } else {
        const int grid = (Bt + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_SOLVE, 32>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, (const float *)nullptr,
               (const float *)nullptr, obj_nrm, (const float *)nullptr, cog, Bt, batch_each, n_c, (float *)nullptr,
               (float *)nullptr, (float *)nullptr, (float *)nullptr, (float *)nullptr, (const int32_t *)nullptr,
               (float *)nullptr, ws, 0, (int32_t *)nullptr);
    }
    GQP_CHECK_LAUNCH("gqp_fc_solve");
    return GQP_OK;
}

extern "C" int gqp_fc_finish(const GqpFcParams *p, const float *cpts, const float *cnrm_hand, const float *obj_dist,
                             const float *obj_nrm, const float *obj_dgrad, const float *cog, int n_obj, int batch_each,
                             int n_c, const void *workspace, int batch_stop, float *e_fc, float *e_dis, float *lam,
                             float *grad_cpts, float *grad_cnrm, int32_t *stop_iter_out, void *stream) {
    FcParamsDev P;
    if (int rc = fc_check("gqp_fc_finish", p, n_obj, batch_each, n_c, P)) return rc;
    const int Bt = n_obj * batch_each;
    if (Bt == 0) return GQP_OK;
    if (!cpts || !cnrm_hand || !obj_dist || !obj_nrm || !cog || !workspace)
        return gqp_set_error(GQP_ERR_ARG, "gqp_fc_finish: null pointer");
    if (!e_fc && !e_dis && !lam && !grad_cpts && !grad_cnrm) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_finish: no output");
    FcWorkspace ws = fc_workspace_view(const_cast<void *>(workspace), Bt);
    // n = 4 n_c <= 32 variables: two candidates per warp (half-warp reductions), else one
    if (n_c <= 8) {
        const int grid = ((Bt + 1) / 2 + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_FINISH, 16>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, cnrm_hand, obj_dist,
               obj_nrm, obj_dgrad, cog, Bt, batch_each, n_c, e_fc, e_dis, lam, grad_cpts, grad_cnrm,
               (const int32_t *)nullptr, (float *)nullptr, ws, batch_stop, stop_iter_out);
    } else {
        const int grid = (Bt + FC_WARPS - 1) / FC_WARPS;
        GQP_LAUNCH((fc_dis_kernel<FC_FINISH, 32>), grid, FC_WARPS * 32, 0, (cudaStream_t)stream, P, cpts, cnrm_hand, obj_dist,
               obj_nrm, obj_dgrad, cog, Bt, batch_each, n_c, e_fc, e_dis, lam, grad_cpts, grad_cnrm,
               (const int32_t *)nullptr, (float *)nullptr, ws, batch_stop, stop_iter_out);
    }
    GQP_CHECK_LAUNCH("gqp_fc_finish");
    return GQP_OK;
}

extern "C" int gqp_fc_stop_iter(const float *qp_trace, int Bt, int max_iter, float eps, int32_t *stop_iter,
                                void *stream) {
    if (!qp_trace || !stop_iter || Bt <= 0 || max_iter < 1) return gqp_set_error(GQP_ERR_ARG, "gqp_fc_stop_iter: bad argument");
    GQP_LAUNCH(fc_stop_iter_kernel, 1, 256, 0, (cudaStream_t)stream, qp_trace, Bt, max_iter, eps, stop_iter);
    GQP_CHECK_LAUNCH("gqp_fc_stop_iter");
    return GQP_OK;
}
