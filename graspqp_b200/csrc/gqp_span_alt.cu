// (f3) The reference's two other force-closure energies + E_dis, forward and analytic backward in ONE launch, a warp per
// candidate (energy_type = dexgrasp | tdg of scripts/fit.py:344-347; the graspqp energy is gqp_fc.cu).
//   dexgrasp : metrics/ops/dexgrasp.py:5-69     E = |sum_i n_i|^2 + tw^2 |sum_i n_i x (p_i - cog)|^2
//   tdg      : metrics/ops/tdg.py:57-239        E = 100 * mean_u (1 - cos(W(u), [u;0])) over P target directions u,
//              W(u) = sum_i rho_i [f_i(u); r_i x f_i(u)], f_i(u) = the point of contact i's normalised friction cone that
//              reaches furthest along u (closed form, three angular regions), r_i = (p_i - cog) / 0.2, rho_i the density weight
//   E_dis    : core/energy.py:25-28 (as in gqp_fc.cu)
// The object normals are constants of the graph (TorchSDF returns them without one, object_model.py:220): only the torque
// rows depend on the contact points, so d E/d p_i = rho_i / 0.2 * sum_u f_i(u) x s(u) with s(u) = dE/dW_tau(u).
//
// Layout: lanes stride the P directions (coalesced reads of the (P,3) table, which stays in L1/L2: 12 KB); the n_c contact
// frames, lever arms and weights of the candidate sit in shared memory (13 floats per contact).  The n_c forces f_i(u) of a
// direction and the per-contact gradient sums are registers: the kernel is instantiated for n_c <= 4 / 8 / 16 / 32 (the last
// one re-evaluates f_i(u) for the gradient instead).
// Bound: FP32 issue — P x n_c cone evaluations of ~60 flop per candidate (0.7 MFLOP at P = 1000, n_c = 12); HBM bytes per
// candidate: the n_c contacts in (150 B each) and their gradients out.
// Measured on a B200 at config 2 (8192 candidates, n_c = 12, P = 1000; profiles/r2p, r2q, r2r): cone forces evaluated twice
// 1.045 ms (25 % of the FP32 FMA probe, 58 % of lane-issue slots); kept in registers between the sum and the gradient
// (150 registers, 12 warps per SM) 0.785 ms; + the direction normalised with one rsqrt instead of sqrt and a division
// 0.678 ms; capping at 128 registers for 16 warps per SM spills and gains nothing (0.688 ms).
#include "gqp_common.cuh"
#include "graspqp_b200.h"

#define SPAN_WARPS 4
#define SPAN_MAX_NC 32

struct SpanContact {
    float a0[3], a1[3], a2[3];  // contact frame, columns (tdg.py:84-107)
    float r[3];                 // (p - cog) / obb_length
    float rho;                  // density weight (tdg.py:129-132), 1 without it
};

__device__ __forceinline__ V3 span_unit(V3 v) {  // tdg.py:37-39: v / clamp(|v|, 1e-12)
    const float n = sqrtf(dot3(v, v));
    return v * (1.0f / fmaxf(n, 1e-12f));
}
// the same through the hardware reciprocal square root (2 ulp): one MUFU instead of a square root and a division
__device__ __forceinline__ V3 span_unit_fast(V3 v) { return v * rsqrtf(fmaxf(dot3(v, v), 1e-24f)); }

// the point of the normalised cone {(1, y, z): |(y,z)| <= mu} (and its scaled copies for directions behind the contact)
// that the reference picks for the direction u: tdg.py:173-195, then rotated to the world frame
__device__ __forceinline__ V3 span_cone_force(const SpanContact &c, V3 u, float mu, float cos_ba) {
    const V3 raw = v3(c.a0[0] * u.x + c.a0[1] * u.y + c.a0[2] * u.z, c.a1[0] * u.x + c.a1[1] * u.y + c.a1[2] * u.z,
                      c.a2[0] * u.x + c.a2[1] * u.y + c.a2[2] * u.z);
    const V3 d = span_unit_fast(raw);
    const float proj = d.x;
    const float pn = sqrtf(d.y * d.y + d.z * d.z);
    float cy, cz, cx = 1.0f;
    if (proj >= cos_ba) {  // region 1: inside the cone — the ray itself
        const float inv = 1.0f / fmaxf(proj, 0.5f * cos_ba);
        cy = d.y * inv;
        cz = d.z * inv;
    } else {               // regions 2, 3: the rim, shrunk for directions pointing away from the contact
        const float inv = mu / fmaxf(pn, 1e-12f);
        cy = d.y * inv;
        cz = d.z * inv;
        if (proj < 0.0f) {
            const float h3 = pn / (pn - 2.0f * mu * proj);
            cx = h3;
            cy *= h3;
            cz *= h3;
        }
    }
    return v3(c.a0[0] * cx + c.a1[0] * cy + c.a2[0] * cz, c.a0[1] * cx + c.a1[1] * cy + c.a2[1] * cz,
              c.a0[2] * cx + c.a1[2] * cy + c.a2[2] * cz);
}

// E_dis and its gradient (energy.py:25-28), lane = contact; returns the candidate's E_dis on every lane and the lane's
// d(w_dis E_dis)/d p through the object distance in gpd
__device__ __forceinline__ float span_e_dis(const GqpSpanParams &P, bool has, bool live, size_t q, size_t qd, V3 n,
                                            const float *__restrict__ cnrm_hand, const float *__restrict__ obj_dist,
                                            const float *__restrict__ obj_dgrad, float *__restrict__ grad_cnrm, V3 &gpd) {
    float edis = 0.f;
    gpd = v3(0.f, 0.f, 0.f);
    if (has) {
        const V3 nh = ld3(cnrm_hand + q);
        const float dd = obj_dist[qd];
        const float ea = expf(1.0f + dot3(n, nh));
        edis = ea * fabsf(dd);
        if (grad_cnrm && live) st3(grad_cnrm + q, n * (P.w_dis * edis));
        const float gd = P.w_dis * ea * (dd > 0.f ? 1.f : (dd < 0.f ? -1.f : 0.f));
        if (obj_dgrad) gpd = ld3(obj_dgrad + q) * gd;
    }
    return warp_sum(edis);
}

// NC: compile-time bound of n_c (4, 8, 16, 32) — the per-contact gradient sums of a lane's directions are NC x 3 registers
// KEEP: the n_c world-frame cone forces of a direction stay in registers between the sum W(u) and the gradient (NC x 3 more
// registers per lane) instead of being evaluated a second time
template <int NC, bool KEEP>
__global__ void __launch_bounds__(SPAN_WARPS * 32)
span_alt_kernel(GqpSpanParams P, const float *__restrict__ cpts, const float *__restrict__ cnrm_hand,
                const float *__restrict__ obj_dist, const float *__restrict__ obj_nrm,
                const float *__restrict__ obj_dgrad, const float *__restrict__ cog, const float *__restrict__ dirs,
                int n_dirs, int Bt, int batch_each, int n_c, float *__restrict__ e_fc, float *__restrict__ e_dis,
                float *__restrict__ grad_cpts, float *__restrict__ grad_cnrm) {
    __shared__ SpanContact sc[SPAN_WARPS][NC];
    __shared__ float sn[SPAN_WARPS][NC][3];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int b = blockIdx.x * SPAN_WARPS + w;
    const bool live = b < Bt;
    const int bb = live ? b : Bt - 1;  // idle warps shadow the last candidate: every lane takes every collective
    const V3 cg = ld3(cog + (size_t)(bb / batch_each) * 3);
    const bool has = lane < n_c;
    const size_t q = ((size_t)bb * n_c + (has ? lane : 0)) * 3;
    const V3 n = ld3(obj_nrm + q), p = ld3(cpts + q) - cg;
    V3 gpd;
    const float edis = span_e_dis(P, has, live, q, (size_t)bb * n_c + (has ? lane : 0), n, cnrm_hand, obj_dist, obj_dgrad,
                                  grad_cnrm, gpd);
    float efc;
    V3 gp = v3(0.f, 0.f, 0.f);
    if (P.kind == GQP_SPAN_DEXGRASP) {
        // contact_normals @ [I; skew(p) tw]: force part sum n, torque part tw * sum n x p (dexgrasp.py:29-36)
        const V3 t = has ? cross3(n, p) * P.torque_weight : v3(0.f, 0.f, 0.f);
        const V3 nn = has ? n : v3(0.f, 0.f, 0.f);
        const V3 sf = v3(warp_sum(nn.x), warp_sum(nn.y), warp_sum(nn.z));
        const V3 stq = v3(warp_sum(t.x), warp_sum(t.y), warp_sum(t.z));
        const float nrm = sqrtf(dot3(sf, sf) + dot3(stq, stq));
        efc = nrm * nrm;
        // d/dp_i of |S|^2, S = tw sum n_i x p_i:  2 tw (S x n_i)
        if (has) gp = cross3(stq, n) * (2.0f * P.torque_weight * P.w_fc);
    } else {
        // ---- contact frames, lever arms, density into shared memory ----
        if (has) {
            sn[w][lane][0] = n.x; sn[w][lane][1] = n.y; sn[w][lane][2] = n.z;
        }
        __syncwarp();
        if (has) {
            SpanContact &c = sc[w][lane];
            const V3 base = fabsf(n.y) > 0.99f ? v3(0.f, 0.f, 1.f) : v3(0.f, 1.f, 0.f);
            V3 a1 = span_unit(base - n * dot3(base, n));
            a1 = span_unit(a1 - n * dot3(a1, n));
            const V3 a2 = cross3(n, a1);
            c.a0[0] = n.x; c.a0[1] = n.y; c.a0[2] = n.z;
            c.a1[0] = a1.x; c.a1[1] = a1.y; c.a1[2] = a1.z;
            c.a2[0] = a2.x; c.a2[1] = a2.y; c.a2[2] = a2.z;
            const float io = 1.0f / P.obb_length;
            c.r[0] = p.x * io; c.r[1] = p.y * io; c.r[2] = p.z * io;
            float rho = 1.0f;
            if (P.enable_density) {
                float s = 0.f;
                for (int j = 0; j < n_c; ++j)
                    s += fmaxf(n.x * sn[w][j][0] + n.y * sn[w][j][1] + n.z * sn[w][j][2], 0.f);
                rho = 1.0f / fmaxf(s, 1e-4f);
            }
            c.rho = rho;
        }
        __syncwarp();
        const float cos_ba = 1.0f / sqrtf(1.0f + P.friction * P.friction);  // cos(atan(mu))
        float esum = 0.f;
        V3 g[NC];
#pragma unroll
        for (int i = 0; i < NC; ++i) g[i] = v3(0.f, 0.f, 0.f);
        for (int u = lane; u < n_dirs; u += 32) {
            const V3 d = ld3(dirs + (size_t)u * 3);
            V3 Wf = v3(0.f, 0.f, 0.f), Wt = v3(0.f, 0.f, 0.f);
            V3 fk[KEEP ? NC : 1];
            if (KEEP) {
#pragma unroll
                for (int i = 0; i < NC; ++i)
                    if (i < n_c) {
                        const SpanContact &c = sc[w][i];
                        const V3 f = span_cone_force(c, d, P.friction, cos_ba) * c.rho;
                        fk[KEEP ? i : 0] = f;
                        Wf = Wf + f;
                        Wt = Wt + cross3(v3(c.r[0], c.r[1], c.r[2]), f);
                    }
            } else {
                for (int i = 0; i < n_c; ++i) {
                    const SpanContact &c = sc[w][i];
                    const V3 f = span_cone_force(c, d, P.friction, cos_ba) * c.rho;
                    Wf = Wf + f;
                    Wt = Wt + cross3(v3(c.r[0], c.r[1], c.r[2]), f);
                }
            }
            const float w2 = dot3(Wf, Wf) + dot3(Wt, Wt);
            const float wn = sqrtf(w2);
            esum += 1.0f - dot3(Wf, d) / fmaxf(wn, 1e-12f);
            // d E/d r_i += rho_i f_i(u) x s(u),  s(u) = (W_f.u) W_tau / |W|^3 (x scale/P below)
            if (grad_cpts && wn > 1e-12f) {
                const V3 s = Wt * (dot3(Wf, d) / (w2 * wn));
#pragma unroll
                for (int i = 0; i < NC; ++i)
                    if (i < n_c) {
                        if (KEEP) {
                            g[i] = g[i] + cross3(fk[KEEP ? i : 0], s);
                        } else {
                            const SpanContact &c = sc[w][i];
                            g[i] = g[i] + cross3(span_cone_force(c, d, P.friction, cos_ba) * c.rho, s);
                        }
                    }
            }
        }
        esum = warp_sum(esum);
        efc = P.scale * esum / (float)n_dirs;
        if (grad_cpts) {
#pragma unroll
            for (int i = 0; i < NC; ++i)
                if (i < n_c) {  // warp-uniform
                    const V3 t = v3(warp_sum(g[i].x), warp_sum(g[i].y), warp_sum(g[i].z));
                    if (lane == i) gp = t * (P.w_fc * P.scale / ((float)n_dirs * P.obb_length));
                }
        }
    }
    if (lane == 0 && live) {
        if (e_fc) e_fc[b] = efc;
        if (e_dis) e_dis[b] = edis;
    }
    if (grad_cpts && has && live) st3(grad_cpts + q, gp + gpd);
}

extern "C" int gqp_span_energy(const GqpSpanParams *p, const float *cpts, const float *cnrm_hand, const float *obj_dist,
                               const float *obj_nrm, const float *obj_dgrad, const float *cog, const float *dirs,
                               int n_dirs, int n_obj, int batch_each, int n_c, float *e_fc, float *e_dis,
                               float *grad_cpts, float *grad_cnrm, void *stream) {
    if (!p || n_obj < 0 || batch_each <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_span_energy: bad argument");
    if (n_c <= 0 || n_c > SPAN_MAX_NC)
        return gqp_set_error(GQP_ERR_LIMIT, "gqp_span_energy: n_c = %d outside 1..%d", n_c, SPAN_MAX_NC);
    if (p->kind != GQP_SPAN_DEXGRASP && p->kind != GQP_SPAN_TDG)
        return gqp_set_error(GQP_ERR_ARG, "gqp_span_energy: kind %d is neither GQP_SPAN_DEXGRASP nor GQP_SPAN_TDG", p->kind);
    if (p->kind == GQP_SPAN_TDG) {
        if (!dirs || n_dirs <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_span_energy: tdg needs the (P,3) direction table");
        if (!(p->friction > 0.f && p->friction <= 1.f) || !(p->obb_length > 0.f))
            return gqp_set_error(GQP_ERR_ARG, "gqp_span_energy: friction must be in (0,1] (tdg.py:66), obb_length > 0");
    }
    const long long Bt = (long long)n_obj * batch_each;
    if (Bt == 0) return GQP_OK;
    if (Bt > 0x7fffffffLL) return gqp_set_error(GQP_ERR_LIMIT, "gqp_span_energy: too many candidates");
    if (!cpts || !cnrm_hand || !obj_dist || !obj_nrm || !cog)
        return gqp_set_error(GQP_ERR_ARG, "gqp_span_energy: null pointer");
    const int grid = (int)((Bt + SPAN_WARPS - 1) / SPAN_WARPS);
#define SPAN_GO(NC, KEEP)                                                                                             \
    GQP_LAUNCH((span_alt_kernel<NC, KEEP>), grid, SPAN_WARPS * 32, 0, (cudaStream_t)stream, *p, cpts, cnrm_hand,      \
               obj_dist, obj_nrm, obj_dgrad, cog, dirs, n_dirs, (int)Bt, batch_each, n_c, e_fc, e_dis, grad_cpts,     \
               grad_cnrm)
    if (n_c <= 4) SPAN_GO(4, true);
    else if (n_c <= 8) SPAN_GO(8, true);
    else if (n_c <= 16) SPAN_GO(16, true);
    else SPAN_GO(32, false);  // 96 + 96 registers of forces and sums would spill: the cone forces are re-evaluated
#undef SPAN_GO
    GQP_CHECK_LAUNCH("gqp_span_energy");
    return GQP_OK;
}
