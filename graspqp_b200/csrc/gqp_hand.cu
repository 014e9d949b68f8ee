// H1/H2/H3 — hand kinematics: forward kinematics over the URDF chain, analytic Jacobian-transpose backward, contact
// candidate gather.  Replaces HandModel.set_parameters / fk / _set_contact_idxs (core/hand_model.py:762-873,
// 1220-1267), pytorch_kinematics' recursive FK (chain.py:146-168, frame.py:98-114, rotation_conversions.py:43-71) and
// the autograd walk through them (scripts/fit.py:440).
//
// Layout: pose (Bt, 9+dof) row-major as in the reference; link_tf (Bt, n_links, 12) world transforms, one candidate's
// links contiguous (the consumers assign one candidate to a warp or block).  Two forms:
//   * the engine's fused kernels (fk_contacts_forward_warp_kernel / fk_contacts_backward_warp_kernel, hands of <= 32
//     frames): a WARP owns a candidate, lane = frame, the tree is swept depth by depth with shuffles (parents sit in
//     other lanes), contact gather and its adjoint in the same launch, transforms never leave registers in between;
//   * the stand-alone gqp_fk_forward / gqp_fk_backward / gqp_contacts_* entry points (HandModel mirror, export, hands
//     with more frames): a thread owns a candidate, tables in shared memory, per-frame transforms in thread-local arrays.
// FK is < 0.1 % of a grasp-step's work; it is bound by nothing but launch latency and the length of the chain.
#include <algorithm>

#include "gqp_common.cuh"
#include "gqp_hand.h"
#include "graspqp_b200.h"

#ifndef FK_THREADS
#define FK_THREADS 128
#endif

__device__ __forceinline__ void load_hand_tables(GqpHandDev *dst, const GqpHandDev *src) {
    const int n = sizeof(GqpHandDev) / 4;
    const int32_t *s = reinterpret_cast<const int32_t *>(src);
    int32_t *d = reinterpret_cast<int32_t *>(dst);
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
}

// utils/transforms.py:5-13 -> roma.special_gramschmidt: columns (x^, y^, x^ x y^)
struct GramSchmidt {
    V3 xh, yh, z;
    float nx, ny, d;  // |x|, |y'|, x^.y
    V3 y;
};
__device__ __forceinline__ GramSchmidt gram_schmidt(V3 x, V3 y) {
    GramSchmidt g;
    g.y = y;
    g.nx = sqrtf(dot3(x, x));
    g.xh = x * (1.0f / g.nx);
    g.d = dot3(g.xh, y);
    V3 yp = y - g.xh * g.d;
    g.ny = sqrtf(dot3(yp, yp));
    g.yh = yp * (1.0f / g.ny);
    g.z = cross3(g.xh, g.yh);
    return g;
}
__device__ __forceinline__ T34 global_tf(const float *pose, GramSchmidt &gs) {
    gs = gram_schmidt(ld3(pose + 3), ld3(pose + 6));
    T34 G;
    G.m[0] = gs.xh.x; G.m[1] = gs.yh.x; G.m[2] = gs.z.x; G.m[3] = pose[0];
    G.m[4] = gs.xh.y; G.m[5] = gs.yh.y; G.m[6] = gs.z.y; G.m[7] = pose[1];
    G.m[8] = gs.xh.z; G.m[9] = gs.yh.z; G.m[10] = gs.z.z; G.m[11] = pose[2];
    return G;
}

// joint transform of frame f for parameter theta (frame.py:98-114, rotation_conversions.py:43-71)
__device__ __forceinline__ T34 joint_tf(const GqpHandDev &H, int f, const float *theta) {
    T34 J = t34_identity();
    int jt = H.frame_jtype[f], d = H.frame_dof[f];
    if (jt == 1) {
        float th = d >= 0 ? theta[d] * H.frame_gain[f] : 0.0f;  // gain 1 unless the joint mimics another (ability hand)
        th = th + 2.0f * ((th > 0.0f ? 1.0f : 0.0f) - 0.5f) * 1e-2f;
        float s, c;
        sincosf(th * 0.5f, &s, &c);
        float r = c, i = s * H.frame_axis[f][0], j = s * H.frame_axis[f][1], k = s * H.frame_axis[f][2];
        float two_s = 2.0f / (r * r + i * i + j * j + k * k);
        J.m[0] = 1 - two_s * (j * j + k * k); J.m[1] = two_s * (i * j - k * r); J.m[2] = two_s * (i * k + j * r);
        J.m[4] = two_s * (i * j + k * r); J.m[5] = 1 - two_s * (i * i + k * k); J.m[6] = two_s * (j * k - i * r);
        J.m[8] = two_s * (i * k - j * r); J.m[9] = two_s * (j * k + i * r); J.m[10] = 1 - two_s * (i * i + j * j);
    } else if (jt == 2 && d >= 0) {
        float th = theta[d] * H.frame_gain[f];
        J.m[3] = th * H.frame_axis[f][0]; J.m[7] = th * H.frame_axis[f][1]; J.m[11] = th * H.frame_axis[f][2];
    }
    return J;
}

// The joint vector the chain is evaluated on: the pose row's joint columns, or, for a hand with a coupling network
// (Robotiq 2F-140: utils/fk.py:23-51, hands/robotiq2.py:9-13), q = [c, W2 tanh(w1 c + b1) + b2] with
// c = clamp(theta_0).  dq (when wanted) = dq/dtheta_0 for the backward; torch.clamp passes the gradient on [lo, hi].
__device__ __forceinline__ const float *joint_vector(const GqpHandDev &H, const float *theta, float *q, float *dq) {
    if (H.mlp_hidden == 0) return theta;
    float c = fminf(fmaxf(theta[0], H.mlp_lo), H.mlp_hi);
    float pass = (theta[0] >= H.mlp_lo && theta[0] <= H.mlp_hi) ? 1.0f : 0.0f;
    q[0] = c;
    if (dq) dq[0] = pass;
    for (int o = 0; o < H.mlp_out; ++o) {
        q[1 + o] = H.mlp_b2[o];
        if (dq) dq[1 + o] = 0.f;
    }
    for (int k = 0; k < H.mlp_hidden; ++k) {
        float t = tanhf(H.mlp_w1[k] * c + H.mlp_b1[k]);
        float dt = (1.0f - t * t) * H.mlp_w1[k] * pass;
        for (int o = 0; o < H.mlp_out; ++o) {
            q[1 + o] += H.mlp_w2[o][k] * t;
            if (dq) dq[1 + o] += H.mlp_w2[o][k] * dt;
        }
    }
    return q;
}

// world transform of every frame BEFORE its link offset (children hang on it): A[f] = A[parent] * (joff * jt);
// B[f] = A[parent] * joff is the joint frame (axis / origin for the backward), returned only when wanted.
template <bool WANT_JOINT_FRAME>
__device__ __forceinline__ void fk_frames(const GqpHandDev &H, const float *theta, const T34 &G, T34 *A, T34 *Bj) {
    for (int f = 0; f < H.n_frames; ++f) {
        int par = H.frame_parent[f];
        T34 joff = t34_load(H.frame_joff[f]);
        T34 J = joint_tf(H, f, theta);
        const T34 &Ap = par < 0 ? G : A[par];
        A[f] = t34_mul(Ap, t34_mul(joff, J));
        if (WANT_JOINT_FRAME) Bj[f] = t34_mul(Ap, joff);
    }
}

__global__ void __launch_bounds__(FK_THREADS)
fk_forward_kernel(const GqpHandDev *__restrict__ hand, const float *__restrict__ pose, int Bt,
                  float *__restrict__ link_tf) {
    __shared__ GqpHandDev H;
    load_hand_tables(&H, hand);
    __syncthreads();
    int b = blockIdx.x * FK_THREADS + threadIdx.x;
    if (b >= Bt) return;
    const int D = 9 + H.n_dofs;
    float prow[9 + GQP_MAX_DOFS];
    for (int i = 0; i < D; ++i) prow[i] = pose[(size_t)b * D + i];
    GramSchmidt gs;
    T34 G = global_tf(prow, gs);
    T34 A[GQP_MAX_FRAMES];
    float q[1 + GQP_MAX_MLP_OUT];
    fk_frames<false>(H, joint_vector(H, prow + 9, q, nullptr), G, A, nullptr);
    for (int l = 0; l < H.n_links; ++l) {
        int f = H.link_frame[l];
        T34 W = t34_mul(A[f], t34_load(H.frame_loff[f]));
        float *o = link_tf + ((size_t)b * H.n_links + l) * 12;
#pragma unroll
        for (int i = 0; i < 12; ++i) o[i] = W.m[i];
    }
}

__global__ void __launch_bounds__(FK_THREADS)
fk_backward_kernel(const GqpHandDev *__restrict__ hand, const float *__restrict__ pose,
                   const float *__restrict__ grad_link_tf, const float *__restrict__ grad_pose_extra, int Bt,
                   float *__restrict__ grad_pose) {
    __shared__ GqpHandDev H;
    load_hand_tables(&H, hand);
    __syncthreads();
    int b = blockIdx.x * FK_THREADS + threadIdx.x;
    if (b >= Bt) return;
    const int D = 9 + H.n_dofs;
    float prow[9 + GQP_MAX_DOFS], gp[9 + GQP_MAX_DOFS];
    for (int i = 0; i < D; ++i) prow[i] = pose[(size_t)b * D + i];
    for (int i = 0; i < 9 + GQP_MAX_DOFS; ++i) gp[i] = 0.f;  // joint slots: one per entry of the joint vector
    GramSchmidt gs;
    T34 G = global_tf(prow, gs);
    T34 A[GQP_MAX_FRAMES], Bj[GQP_MAX_FRAMES];
    float q[1 + GQP_MAX_MLP_OUT], dq[1 + GQP_MAX_MLP_OUT];
    fk_frames<true>(H, joint_vector(H, prow + 9, q, dq), G, A, Bj);
    // wrench (torque about the world origin, force) that the loss exerts on every frame
    float accM[GQP_MAX_FRAMES][3], accF[GQP_MAX_FRAMES][3];
    for (int f = 0; f < H.n_frames; ++f)
        for (int k = 0; k < 3; ++k) accM[f][k] = accF[f][k] = 0.f;
    float S[9];  // sum_l gW.R W.R^T + gW.t (W.t - t_g)^T
    for (int i = 0; i < 9; ++i) S[i] = 0.f;
    V3 tg = V3{prow[0], prow[1], prow[2]};
    for (int l = 0; l < H.n_links; ++l) {
        int f = H.link_frame[l];
        T34 W = t34_mul(A[f], t34_load(H.frame_loff[f]));
        T34 g = t34_load(grad_link_tf + ((size_t)b * H.n_links + l) * 12);
        V3 gt = V3{g.m[3], g.m[7], g.m[11]}, wt = V3{W.m[3], W.m[7], W.m[11]};
        V3 M = cross3(wt, gt);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            V3 rk = V3{W.m[k], W.m[4 + k], W.m[8 + k]}, gk = V3{g.m[k], g.m[4 + k], g.m[8 + k]};
            M = M + cross3(rk, gk);
        }
        accM[f][0] += M.x; accM[f][1] += M.y; accM[f][2] += M.z;
        accF[f][0] += gt.x; accF[f][1] += gt.y; accF[f][2] += gt.z;
        V3 dt = wt - tg;
        float dtv[3] = {dt.x, dt.y, dt.z}, gtv[3] = {gt.x, gt.y, gt.z};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                S[3 * i + j] += g.m[4 * i] * W.m[4 * j] + g.m[4 * i + 1] * W.m[4 * j + 1] +
                                g.m[4 * i + 2] * W.m[4 * j + 2] + gtv[i] * dtv[j];
    }
    V3 Ftot = V3{0, 0, 0};
    for (int f = H.n_frames - 1; f >= 0; --f) {
        V3 M = V3{accM[f][0], accM[f][1], accM[f][2]}, F = V3{accF[f][0], accF[f][1], accF[f][2]};
        int jt = H.frame_jtype[f], d = H.frame_dof[f];
        if (jt != 0 && d >= 0) {
            V3 a = t34_vec(Bj[f], V3{H.frame_axis[f][0], H.frame_axis[f][1], H.frame_axis[f][2]});
            if (jt == 1) {
                V3 o = V3{Bj[f].m[3], Bj[f].m[7], Bj[f].m[11]};
                gp[9 + d] += H.frame_gain[f] * dot3(a, M - cross3(o, F));
            } else {
                gp[9 + d] += H.frame_gain[f] * dot3(a, F);
            }
        }
        int par = H.frame_parent[f];
        if (par >= 0) {
            accM[par][0] += M.x; accM[par][1] += M.y; accM[par][2] += M.z;
            accF[par][0] += F.x; accF[par][1] += F.y; accF[par][2] += F.z;
        } else {
            Ftot = Ftot + F;
        }
    }
    if (H.mlp_hidden > 0) {  // chain rule through the coupling network: d/dtheta_0 = sum_j d/dq_j dq_j/dtheta_0
        float g = 0.f;
        for (int j = 0; j <= H.mlp_out; ++j) g += gp[9 + j] * dq[j];
        gp[9] = g;
    }
    gp[0] = Ftot.x; gp[1] = Ftot.y; gp[2] = Ftot.z;
    // dL/dR_g = S R_g, columns -> (x^, y^, z); then back through the Gram-Schmidt
    float Rg[9] = {G.m[0], G.m[1], G.m[2], G.m[4], G.m[5], G.m[6], G.m[8], G.m[9], G.m[10]};
    float gR[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) gR[3 * i + j] = S[3 * i] * Rg[j] + S[3 * i + 1] * Rg[3 + j] + S[3 * i + 2] * Rg[6 + j];
    V3 gxh = V3{gR[0], gR[3], gR[6]}, gyh = V3{gR[1], gR[4], gR[7]}, gz = V3{gR[2], gR[5], gR[8]};
    gxh = gxh + cross3(gs.yh, gz);
    gyh = gyh + cross3(gz, gs.xh);
    V3 gyp = (gyh - gs.yh * dot3(gs.yh, gyh)) * (1.0f / gs.ny);
    V3 gy = gyp;
    float gd = -dot3(gs.xh, gyp);
    gxh = gxh - gyp * gs.d;
    gxh = gxh + gs.y * gd;
    gy = gy + gs.xh * gd;
    V3 gx = (gxh - gs.xh * dot3(gs.xh, gxh)) * (1.0f / gs.nx);
    gp[3] = gx.x; gp[4] = gx.y; gp[5] = gx.z;
    gp[6] = gy.x; gp[7] = gy.y; gp[8] = gy.z;
    for (int i = 0; i < D; ++i)
        grad_pose[(size_t)b * D + i] = gp[i] + (grad_pose_extra ? grad_pose_extra[(size_t)b * D + i] : 0.f);
}

// ---- H3: contact candidates -> world contact points / normals (hand_model.py:787-831, 1220-1267) -------------------
// forward: a thread per (candidate, contact).  The reference transforms all C candidates and gathers n_c of them.
__global__ void contacts_forward_kernel(const float *__restrict__ cand, const float *__restrict__ cand_normal,
                                        const int32_t *__restrict__ cand_link, int n_cand, int n_links,
                                        const float *__restrict__ link_tf, const int32_t *__restrict__ idx, int n,
                                        int n_c, float *__restrict__ cpts, float *__restrict__ cnrm) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int b = i / n_c;
    int k = idx[i];
    k = k < 0 ? 0 : (k >= n_cand ? n_cand - 1 : k);
    int l = cand_link[k];
    T34 W = t34_load(link_tf + ((size_t)b * n_links + l) * 12);
    st3(cpts + 3 * (size_t)i, t34_point(W, ld3(cand + 3 * k)));
    st3(cnrm + 3 * (size_t)i, t34_vec(W, ld3(cand_normal + 3 * k)));
}
// backward: a thread per candidate walks its n_c contacts; accumulates into grad_link_tf with atomics (the penetration
// and self-penetration kernels add to the same buffer, possibly concurrently from another stream).
__global__ void contacts_backward_kernel(const float *__restrict__ cand, const float *__restrict__ cand_normal,
                                         const int32_t *__restrict__ cand_link, int n_cand, int n_links,
                                         const int32_t *__restrict__ idx, const float *__restrict__ grad_cpts,
                                         const float *__restrict__ grad_cnrm, int Bt, int n_c,
                                         float *__restrict__ grad_link_tf) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    for (int c = 0; c < n_c; ++c) {
        size_t i = (size_t)b * n_c + c;
        int k = idx[i];
        k = k < 0 ? 0 : (k >= n_cand ? n_cand - 1 : k);
        int l = cand_link[k];
        V3 p = ld3(cand + 3 * k), nn = ld3(cand_normal + 3 * k);
        V3 gpnt = grad_cpts ? ld3(grad_cpts + 3 * i) : V3{0, 0, 0};
        V3 gn = grad_cnrm ? ld3(grad_cnrm + 3 * i) : V3{0, 0, 0};
        float *g = grad_link_tf + ((size_t)b * n_links + l) * 12;
        float gpv[3] = {gpnt.x, gpnt.y, gpnt.z}, gnv[3] = {gn.x, gn.y, gn.z};
        float pv[3] = {p.x, p.y, p.z}, nv[3] = {nn.x, nn.y, nn.z};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
            for (int q = 0; q < 3; ++q) atomicAdd(g + 4 * r + q, gpv[r] * pv[q] + gnv[r] * nv[q]);
            atomicAdd(g + 4 * r + 3, gpv[r]);
        }
    }
}

// ---- warp per candidate (north_star stage 1): lane = frame of the kinematic tree ------------------------------------------
// The thread-per-candidate kernels above walk 19-27 frames serially with ~2 KB of thread-local transforms: 0.03 + 0.05 ms
// of pure latency on the critical chain at 8192 candidates, on 64 of 148 SMs.  Here a warp owns a candidate: every lane
// builds its frame's local transform, the tree is swept level by level (a lane multiplies its parent's world transform,
// read from shared memory, by its own local one), and the contact gather / scatter rides in the same launch.  The
// backward sweeps the levels upwards; a parent sums its children's wrenches in child order and every reduction has a
// fixed order, so the result is run-to-run reproducible (no atomics).  Requires n_frames <= 32 and n_links <= 32.
#define FKW_WARPS 4
struct FkwLane {
    T34 A;    // world transform of the frame before its link offset
    T34 Bj;   // world joint frame A[parent] * joff (axis / origin of the joint for the backward)
    GramSchmidt gs;
    T34 G;
};
// the sweep shared by both kernels; sA: this warp's 32 x 12 floats
template <bool WANT_JOINT_FRAME>
__device__ __forceinline__ void fkw_sweep(const GqpHandDev &H, const float *prow, int f, bool active, float *sA,
                                          FkwLane &o, float *dq) {
    float p9[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) p9[i] = prow[i];
    o.G = global_tf(p9, o.gs);
    float q[1 + GQP_MAX_MLP_OUT];
    const float *theta = joint_vector(H, prow + 9, q, dq);
    T34 joff = t34_identity(), Lc = t34_identity();
    if (active) {
        joff = t34_load(H.frame_joff[f]);
        Lc = t34_mul(joff, joint_tf(H, f, theta));
    }
    const int depth = active ? H.frame_depth[f] : -1, par = active ? H.frame_parent[f] : -1;
    for (int d = 0; d <= H.max_depth; ++d) {
        if (depth == d) {
            T34 Ap = o.G;
            if (par >= 0) Ap = t34_load(sA + 12 * par);
            o.A = t34_mul(Ap, Lc);
            if (WANT_JOINT_FRAME) o.Bj = t34_mul(Ap, joff);
#pragma unroll
            for (int i = 0; i < 12; ++i) sA[12 * f + i] = o.A.m[i];
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(FKW_WARPS * 32)
fk_contacts_forward_warp_kernel(const GqpHandDev *__restrict__ hand, const float *__restrict__ cand,
                                const float *__restrict__ cand_normal, const int32_t *__restrict__ cand_link,
                                const float *__restrict__ pose, const int32_t *__restrict__ idx, int Bt, int n_c,
                                float *__restrict__ link_tf, float *__restrict__ cpts, float *__restrict__ cnrm) {
    __shared__ float sA[FKW_WARPS][32 * 12];
    __shared__ float sW[FKW_WARPS][GQP_MAX_LINKS * 12];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x * FKW_WARPS + warp;
    if (b >= Bt) return;
    const GqpHandDev &H = *hand;
    const int D = 9 + H.n_dofs, L = H.n_links;
    const bool active = lane < H.n_frames;
    FkwLane st;
    fkw_sweep<false>(H, pose + (size_t)b * D, lane, active, sA[warp], st, nullptr);
    if (active && H.frame_link[lane] >= 0) {
        const T34 W = t34_mul(st.A, t34_load(H.frame_loff[lane]));
        float *o = sW[warp] + 12 * H.frame_link[lane];
#pragma unroll
        for (int i = 0; i < 12; ++i) o[i] = W.m[i];
    }
    __syncwarp();
    for (int i = lane; i < L * 12; i += 32) link_tf[(size_t)b * L * 12 + i] = sW[warp][i];
    if (idx)
        for (int c = lane; c < n_c; c += 32) {
            const size_t i = (size_t)b * n_c + c;
            int k = idx[i];
            k = k < 0 ? 0 : (k >= H.n_cand ? H.n_cand - 1 : k);
            const T34 W = t34_load(sW[warp] + 12 * cand_link[k]);
            st3(cpts + 3 * i, t34_point(W, ld3(cand + 3 * k)));
            st3(cnrm + 3 * i, t34_vec(W, ld3(cand_normal + 3 * k)));
        }
}

__global__ void __launch_bounds__(FKW_WARPS * 32)
fk_contacts_backward_warp_kernel(const GqpHandDev *__restrict__ hand, const float *__restrict__ cand,
                                 const float *__restrict__ cand_normal, const int32_t *__restrict__ cand_link,
                                 const float *__restrict__ pose, const int32_t *__restrict__ idx,
                                 const float *__restrict__ grad_cpts, const float *__restrict__ grad_cnrm,
                                 const float *__restrict__ grad_link_tf, const float *__restrict__ grad_pose_extra,
                                 int Bt, int n_c, float *__restrict__ grad_pose) {
    __shared__ float sA[FKW_WARPS][32 * 12];
    __shared__ float sG[FKW_WARPS][GQP_MAX_LINKS * 12];  // d loss / d link_tf incl. the contacts
    __shared__ float sMF[FKW_WARPS][32 * 6];             // wrench of every frame's subtree
    __shared__ float sJ[FKW_WARPS][32];                  // joint-gradient contribution of every frame
    __shared__ int sCk[FKW_WARPS][32];                   // contact -> candidate id / link (n_c <= 32 per round)
    __shared__ int sCl[FKW_WARPS][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x * FKW_WARPS + warp;
    if (b >= Bt) return;
    const GqpHandDev &H = *hand;
    const int D = 9 + H.n_dofs, L = H.n_links, f = lane;
    const bool active = lane < H.n_frames;
    const float *prow = pose + (size_t)b * D;
    FkwLane st;
    float dq[1 + GQP_MAX_MLP_OUT];
    fkw_sweep<true>(H, prow, lane, active, sA[warp], st, dq);
    // ---- d loss / d link_tf of link `lane`: what the other stages left in global memory + this candidate's contacts, in
    //      contact order (contacts_backward_kernel's formula)
    float g[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) g[i] = (lane < L && grad_link_tf) ? grad_link_tf[((size_t)b * L + lane) * 12 + i] : 0.f;
    for (int c0 = 0; idx && c0 < n_c; c0 += 32) {
        const int c = c0 + lane;
        if (c < n_c) {
            int k = idx[(size_t)b * n_c + c];
            k = k < 0 ? 0 : (k >= H.n_cand ? H.n_cand - 1 : k);
            sCk[warp][lane] = k;
            sCl[warp][lane] = cand_link[k];
        }
        __syncwarp();
        const int nc = min(32, n_c - c0);
        for (int j = 0; j < nc; ++j) {
            if (sCl[warp][j] != lane) continue;
            const int k = sCk[warp][j];
            const size_t i = (size_t)b * n_c + c0 + j;
            const V3 p = ld3(cand + 3 * k), nn = ld3(cand_normal + 3 * k);
            const V3 gp = grad_cpts ? ld3(grad_cpts + 3 * i) : V3{0, 0, 0}, gn = grad_cnrm ? ld3(grad_cnrm + 3 * i) : V3{0, 0, 0};
            const float gpv[3] = {gp.x, gp.y, gp.z}, gnv[3] = {gn.x, gn.y, gn.z}, pv[3] = {p.x, p.y, p.z},
                        nv[3] = {nn.x, nn.y, nn.z};
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int q2 = 0; q2 < 3; ++q2) g[4 * r + q2] += gpv[r] * pv[q2] + gnv[r] * nv[q2];
                g[4 * r + 3] += gpv[r];
            }
        }
        __syncwarp();
    }
    if (lane < L) {
#pragma unroll
        for (int i = 0; i < 12; ++i) sG[warp][12 * lane + i] = g[i];
    }
    __syncwarp();
    // ---- wrench (torque about the world origin, force) the loss exerts on this lane's frame through its link ----
    float M[3] = {0.f, 0.f, 0.f}, F[3] = {0.f, 0.f, 0.f}, S[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) S[i] = 0.f;
    const V3 tg = V3{prow[0], prow[1], prow[2]};
    if (active && H.frame_link[f] >= 0) {
        const T34 W = t34_mul(st.A, t34_load(H.frame_loff[f]));
        const T34 gl = t34_load(sG[warp] + 12 * H.frame_link[f]);
        const V3 gt = V3{gl.m[3], gl.m[7], gl.m[11]}, wt = V3{W.m[3], W.m[7], W.m[11]};
        V3 Mv = cross3(wt, gt);
#pragma unroll
        for (int k = 0; k < 3; ++k)
            Mv = Mv + cross3(V3{W.m[k], W.m[4 + k], W.m[8 + k]}, V3{gl.m[k], gl.m[4 + k], gl.m[8 + k]});
        M[0] = Mv.x; M[1] = Mv.y; M[2] = Mv.z;
        F[0] = gt.x; F[1] = gt.y; F[2] = gt.z;
        const V3 dt = wt - tg;
        const float dtv[3] = {dt.x, dt.y, dt.z};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                S[3 * i + j] = gl.m[4 * i] * W.m[4 * j] + gl.m[4 * i + 1] * W.m[4 * j + 1] + gl.m[4 * i + 2] * W.m[4 * j + 2] +
                               F[i] * dtv[j];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) S[i] = warp_sum(S[i]);
    // ---- upward sweep: a frame's wrench = its own + its children's (child order), then its joint's gradient ----
    const int depth = active ? H.frame_depth[f] : -1;
    float jg = 0.f;
    for (int d = H.max_depth; d >= 0; --d) {
        if (depth == d) {
            const int nch = min(H.frame_nchild[f], GQP_MAX_CHILDREN);
            for (int k = 0; k < nch; ++k) {
                const float *ch = sMF[warp] + 6 * H.frame_child[f][k];
                M[0] += ch[0]; M[1] += ch[1]; M[2] += ch[2];
                F[0] += ch[3]; F[1] += ch[4]; F[2] += ch[5];
            }
            float *me = sMF[warp] + 6 * f;
            me[0] = M[0]; me[1] = M[1]; me[2] = M[2]; me[3] = F[0]; me[4] = F[1]; me[5] = F[2];
            const int jt = H.frame_jtype[f], dd = H.frame_dof[f];
            if (jt != 0 && dd >= 0) {
                const V3 a = t34_vec(st.Bj, V3{H.frame_axis[f][0], H.frame_axis[f][1], H.frame_axis[f][2]});
                const V3 Mv = V3{M[0], M[1], M[2]}, Fv = V3{F[0], F[1], F[2]};
                if (jt == 1) jg = dot3(a, Mv - cross3(V3{st.Bj.m[3], st.Bj.m[7], st.Bj.m[11]}, Fv));
                else jg = dot3(a, Fv);
                jg *= H.frame_gain[f];
            }
        }
        __syncwarp();
    }
    sJ[warp][lane] = jg;
    __syncwarp();
    // ---- joint gradients: entry j of the joint vector sums the frames it drives, in frame order ----
    float *out = grad_pose + (size_t)b * D;
    const int n_q = H.mlp_hidden > 0 ? 1 + H.mlp_out : H.n_dofs;
    float gq = 0.f;
    if (lane < n_q)
        for (int ff = 0; ff < H.n_frames; ++ff)
            if (H.frame_jtype[ff] != 0 && H.frame_dof[ff] == lane) gq += sJ[warp][ff];
    if (H.mlp_hidden > 0) {  // chain rule through the coupling network: d/dtheta_0 = sum_j d/dq_j dq_j/dtheta_0
        float t = lane < n_q ? gq * dq[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) out[9] = t + (grad_pose_extra ? grad_pose_extra[(size_t)b * D + 9] : 0.f);
    } else if (lane < n_q) {
        out[9 + lane] = gq + (grad_pose_extra ? grad_pose_extra[(size_t)b * D + 9 + lane] : 0.f);
    }
    if (lane == 0) {
        V3 Ftot = V3{0, 0, 0};
        for (int ff = 0; ff < H.n_frames; ++ff)
            if (H.frame_parent[ff] < 0) Ftot = Ftot + V3{sMF[warp][6 * ff + 3], sMF[warp][6 * ff + 4], sMF[warp][6 * ff + 5]};
        float gp[9];
        gp[0] = Ftot.x; gp[1] = Ftot.y; gp[2] = Ftot.z;
        // dL/dR_g = S R_g, columns -> (x^, y^, z); then back through the Gram-Schmidt (as fk_backward_kernel)
        const T34 &G = st.G;
        const GramSchmidt &gs = st.gs;
        const float Rg[9] = {G.m[0], G.m[1], G.m[2], G.m[4], G.m[5], G.m[6], G.m[8], G.m[9], G.m[10]};
        float gR[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) gR[3 * i + j] = S[3 * i] * Rg[j] + S[3 * i + 1] * Rg[3 + j] + S[3 * i + 2] * Rg[6 + j];
        V3 gxh = V3{gR[0], gR[3], gR[6]}, gyh = V3{gR[1], gR[4], gR[7]}, gz = V3{gR[2], gR[5], gR[8]};
        gxh = gxh + cross3(gs.yh, gz);
        gyh = gyh + cross3(gz, gs.xh);
        const V3 gyp = (gyh - gs.yh * dot3(gs.yh, gyh)) * (1.0f / gs.ny);
        V3 gy = gyp;
        const float gd = -dot3(gs.xh, gyp);
        gxh = gxh - gyp * gs.d;
        gxh = gxh + gs.y * gd;
        gy = gy + gs.xh * gd;
        const V3 gx = (gxh - gs.xh * dot3(gs.xh, gxh)) * (1.0f / gs.nx);
        gp[3] = gx.x; gp[4] = gx.y; gp[5] = gx.z;
        gp[6] = gy.x; gp[7] = gy.y; gp[8] = gy.z;
        for (int i = 0; i < 9; ++i) out[i] = gp[i] + (grad_pose_extra ? grad_pose_extra[(size_t)b * D + i] : 0.f);
    }
}

// ---- (f3) optional energy terms E_prior / E_wall (core/energy.py:68-76, weights --w_prior / --w_wall of scripts/fit.py) ----
// E_prior = 1 - (R_g a) . (0,0,-1) = 1 + (R_g a)_z with a = HandModel.grasp_axis: the grasp axis should point down.
// E_wall  = sum over the hand surface samples of |min(z, 0)|: nothing below the table plane z = 0 (hand_model.py:1042-1071).
// A warp per candidate: lanes stride the surface samples; gradients: d E_wall / d link_tf row 2 by atomics (the buffer the
// penetration kernels add to), d E_prior / d pose[3:9] through the Gram-Schmidt (lane 0), ADDED to grad_pose.
__device__ __forceinline__ void gs_backward_row2(const GramSchmidt &gs, float gx_z, float gy_z, float gz_z, V3 &gx, V3 &gy) {
    // upstream gradient w.r.t. the THIRD ROW of R = [x^ y^ z]: (gx_z, gy_z, gz_z); same chain as fk_backward_kernel
    V3 gxh = V3{0.f, 0.f, gx_z}, gyh = V3{0.f, 0.f, gy_z}, gz = V3{0.f, 0.f, gz_z};
    gxh = gxh + cross3(gs.yh, gz);
    gyh = gyh + cross3(gz, gs.xh);
    const V3 gyp = (gyh - gs.yh * dot3(gs.yh, gyh)) * (1.0f / gs.ny);
    gy = gyp;
    const float gd = -dot3(gs.xh, gyp);
    gxh = gxh - gyp * gs.d;
    gxh = gxh + gs.y * gd;
    gy = gy + gs.xh * gd;
    gx = (gxh - gs.xh * dot3(gs.xh, gxh)) * (1.0f / gs.nx);
}

__global__ void __launch_bounds__(128)
prior_wall_kernel(int n_links, const float *__restrict__ link_tf, const float *__restrict__ pose, int Bt, int D, float ax,
                  float ay, float az, const float *__restrict__ surf, const int32_t *__restrict__ surf_link, int n_surf,
                  float w_prior, float w_wall, float *__restrict__ e_prior, float *__restrict__ e_wall,
                  float *__restrict__ grad_link_tf, float *__restrict__ grad_pose) {
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const bool live = b < Bt;
    const int bb = live ? b : Bt - 1;
    float ew = 0.f;
    if (e_wall || (grad_link_tf && w_wall != 0.f)) {
        for (int i = lane; i < n_surf; i += 32) {
            const int l = surf_link[i];
            const float *W = link_tf + ((size_t)bb * n_links + l) * 12;
            const V3 p = ld3(surf + 3 * (size_t)i);
            const float z = W[8] * p.x + W[9] * p.y + W[10] * p.z + W[11];
            if (z < 0.f) {
                ew -= z;
                if (grad_link_tf && live && w_wall != 0.f) {  // d|z|/dz = -1
                    float *g = grad_link_tf + ((size_t)b * n_links + l) * 12;
                    atomicAdd(g + 8, -w_wall * p.x);
                    atomicAdd(g + 9, -w_wall * p.y);
                    atomicAdd(g + 10, -w_wall * p.z);
                    atomicAdd(g + 11, -w_wall);
                }
            }
        }
        ew = warp_sum(ew);
    }
    if (lane == 0 && live) {
        if (e_wall) e_wall[b] = ew;
        const float *pr = pose + (size_t)b * D;
        const GramSchmidt gs = gram_schmidt(ld3(pr + 3), ld3(pr + 6));
        if (e_prior) e_prior[b] = 1.0f + (gs.xh.z * ax + gs.yh.z * ay + gs.z.z * az);
        if (grad_pose && w_prior != 0.f) {
            V3 gx, gy;
            gs_backward_row2(gs, w_prior * ax, w_prior * ay, w_prior * az, gx, gy);
            float *g = grad_pose + (size_t)b * D;
            g[3] += gx.x; g[4] += gx.y; g[5] += gx.z;
            g[6] += gy.x; g[7] += gy.y; g[8] += gy.z;
        }
    }
}

extern "C" int gqp_prior_wall(const GqpHand *h, const float *link_tf, const float *pose, int Bt,
                              const float *grasp_axis_host, const float *surf, const int32_t *surf_link, int n_surf,
                              float w_prior, float w_wall, float *e_prior, float *e_wall, float *grad_link_tf,
                              float *grad_pose, void *stream) {
    if (!h || Bt < 0 || n_surf < 0 || !grasp_axis_host) return gqp_set_error(GQP_ERR_ARG, "gqp_prior_wall: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!link_tf || !pose || (n_surf > 0 && (!surf || !surf_link)))
        return gqp_set_error(GQP_ERR_ARG, "gqp_prior_wall: null pointer");
    GQP_LAUNCH(prior_wall_kernel, (Bt + 3) / 4, 128, 0, (cudaStream_t)stream, h->host.n_links, link_tf, pose, Bt,
               9 + h->host.n_dofs, grasp_axis_host[0], grasp_axis_host[1], grasp_axis_host[2], surf, surf_link, n_surf,
               w_prior, w_wall, e_prior, e_wall, grad_link_tf, grad_pose);
    GQP_CHECK_LAUNCH("gqp_prior_wall");
    return GQP_OK;
}

// ---- host side ------------------------------------------------------------------------------------------------------
extern "C" int gqp_hand_create(const GqpHandDesc *d, GqpHand **out) {
    if (!d || !out) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_create: null argument");
    if (d->n_frames <= 0 || d->n_frames > GQP_MAX_FRAMES || d->n_links <= 0 || d->n_links > GQP_MAX_LINKS ||
        d->n_dofs < 0 || d->n_dofs > GQP_MAX_DOFS || d->n_cand <= 0 || d->n_spheres < 0)
        return gqp_set_error(GQP_ERR_LIMIT, "gqp_hand_create: sizes out of range (frames %d/%d links %d/%d dofs %d/%d)",
                             d->n_frames, GQP_MAX_FRAMES, d->n_links, GQP_MAX_LINKS, d->n_dofs, GQP_MAX_DOFS);
    if (d->mlp_hidden < 0 || d->mlp_hidden > GQP_MAX_MLP_HIDDEN || d->mlp_out < 0 || d->mlp_out > GQP_MAX_MLP_OUT)
        return gqp_set_error(GQP_ERR_LIMIT, "gqp_hand_create: coupling network %d x %d exceeds %d x %d", d->mlp_hidden,
                             d->mlp_out, GQP_MAX_MLP_HIDDEN, GQP_MAX_MLP_OUT);
    if (d->mlp_hidden > 0 && (d->n_dofs != 1 || !d->mlp_w1_host || !d->mlp_b1_host || !d->mlp_w2_host ||
                              !d->mlp_b2_host))
        return gqp_set_error(GQP_ERR_ARG, "gqp_hand_create: a coupling network needs n_dofs == 1 and all four tables");
    const int n_q = d->mlp_hidden > 0 ? 1 + d->mlp_out : d->n_dofs;  // entries of the joint vector frame_dof indexes
    GqpHand *h = new GqpHand();
    GqpHandDev &H = h->host;
    std::memset(&H, 0, sizeof(H));
    if (d->mlp_hidden > 0) {
        H.mlp_hidden = d->mlp_hidden; H.mlp_out = d->mlp_out; H.mlp_lo = d->mlp_lo; H.mlp_hi = d->mlp_hi;
        for (int k = 0; k < d->mlp_hidden; ++k) {
            H.mlp_w1[k] = d->mlp_w1_host[k];
            H.mlp_b1[k] = d->mlp_b1_host[k];
            for (int o = 0; o < d->mlp_out; ++o) H.mlp_w2[o][k] = d->mlp_w2_host[o * d->mlp_hidden + k];
        }
        for (int o = 0; o < d->mlp_out; ++o) H.mlp_b2[o] = d->mlp_b2_host[o];
    }
    H.n_frames = d->n_frames; H.n_dofs = d->n_dofs; H.n_links = d->n_links; H.n_cand = d->n_cand;
    H.n_spheres = d->n_spheres;
    for (int f = 0; f < d->n_frames; ++f) {
        H.frame_parent[f] = d->frame_parent_host[f];
        H.frame_jtype[f] = d->frame_jtype_host[f];
        H.frame_dof[f] = d->frame_dof_host[f];
        H.frame_gain[f] = d->frame_gain_host ? d->frame_gain_host[f] : 1.0f;
        H.frame_link[f] = -1;
        if (H.frame_parent[f] >= f || H.frame_dof[f] >= n_q) {
            delete h;
            return gqp_set_error(GQP_ERR_ARG, "gqp_hand_create: frame %d: parent must precede it, dof < %d", f, n_q);
        }
        for (int k = 0; k < 3; ++k) H.frame_axis[f][k] = d->frame_axis_host[3 * f + k];
        for (int k = 0; k < 12; ++k) {
            H.frame_joff[f][k] = d->frame_joff_host[12 * f + k];
            H.frame_loff[f][k] = d->frame_loff_host[12 * f + k];
        }
    }
    for (int f = 0; f < d->n_frames; ++f) {  // tree shape (parents precede their children)
        const int par = H.frame_parent[f];
        H.frame_depth[f] = par < 0 ? 0 : H.frame_depth[par] + 1;
        H.max_depth = std::max(H.max_depth, H.frame_depth[f]);
        if (par >= 0) {
            if (H.frame_nchild[par] < GQP_MAX_CHILDREN) H.frame_child[par][H.frame_nchild[par]] = f;
            H.frame_nchild[par]++;
            H.max_children = std::max(H.max_children, H.frame_nchild[par]);
        }
    }
    for (int l = 0; l < d->n_links; ++l) {
        H.link_frame[l] = d->link_frame_host[l];
        H.frame_link[H.link_frame[l]] = l;
    }
    for (int j = 0; j < d->n_dofs; ++j) {
        H.joints_lower[j] = d->joints_lower_host[j];
        H.joints_upper[j] = d->joints_upper_host[j];
    }
    for (int k = 0; k < d->n_cand; ++k)
        if (d->cand_link_host[k] < 0 || d->cand_link_host[k] >= d->n_links) {
            delete h;
            return gqp_set_error(GQP_ERR_ARG, "gqp_hand_create: cand_link[%d] out of range", k);
        }
    h->d_dev = (GqpHandDev *)gqp_dev_alloc_upload(&H, sizeof(H));
    h->d_cand = (float *)gqp_dev_alloc_upload(d->cand_host, sizeof(float) * 3 * d->n_cand);
    h->d_cand_normal = (float *)gqp_dev_alloc_upload(d->cand_normal_host, sizeof(float) * 3 * d->n_cand);
    h->d_cand_link = (int32_t *)gqp_dev_alloc_upload(d->cand_link_host, sizeof(int32_t) * d->n_cand);
    h->d_spheres = (float *)gqp_dev_alloc_upload(d->spheres_host, sizeof(float) * 4 * d->n_spheres);
    h->d_sphere_link = (int32_t *)gqp_dev_alloc_upload(d->sphere_link_host, sizeof(int32_t) * d->n_spheres);
    h->face_offset.assign(d->face_offset_host, d->face_offset_host + d->n_links + 1);
    if (!h->d_dev || !h->d_cand || !h->d_cand_normal || !h->d_cand_link || !h->d_spheres || !h->d_sphere_link) {
        gqp_hand_destroy(h);
        return gqp_set_error(GQP_ERR_CUDA, "gqp_hand_create: device allocation failed");
    }
    int rc = gqp_pen_tables_create(h, d->face_verts_host);
    if (rc != GQP_OK) {
        gqp_hand_destroy(h);
        return rc;
    }
    *out = h;
    return GQP_OK;
}

extern "C" void gqp_hand_destroy(GqpHand *h) {
    if (!h) return;
    gqp_pen_tables_destroy(h);
    gqp_dev_free(h->d_dev); gqp_dev_free(h->d_cand); gqp_dev_free(h->d_cand_normal); gqp_dev_free(h->d_cand_link);
    gqp_dev_free(h->d_spheres); gqp_dev_free(h->d_sphere_link);
    delete h;
}

extern "C" int gqp_fk_forward(const GqpHand *h, const float *pose, int Bt, float *link_tf, void *stream) {
    if (!h || Bt < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_forward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !link_tf) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_forward: null pointer");
    GQP_LAUNCH(fk_forward_kernel, (Bt + FK_THREADS - 1) / FK_THREADS, FK_THREADS, 0, (cudaStream_t)stream, h->d_dev,
               pose, Bt, link_tf);
    GQP_CHECK_LAUNCH("gqp_fk_forward");
    return GQP_OK;
}

extern "C" int gqp_fk_backward(const GqpHand *h, const float *pose, const float *grad_link_tf,
                               const float *grad_pose_extra, int Bt, float *grad_pose, void *stream) {
    if (!h || Bt < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_backward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !grad_link_tf || !grad_pose) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_backward: null pointer");
    GQP_LAUNCH(fk_backward_kernel, (Bt + FK_THREADS - 1) / FK_THREADS, FK_THREADS, 0, (cudaStream_t)stream, h->d_dev,
               pose, grad_link_tf, grad_pose_extra, Bt, grad_pose);
    GQP_CHECK_LAUNCH("gqp_fk_backward");
    return GQP_OK;
}

extern "C" int gqp_contacts_forward(const GqpHand *h, const float *link_tf, const int32_t *idx, int Bt, int n_c,
                                    float *cpts, float *cnrm, void *stream) {
    if (!h || Bt < 0 || n_c <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_contacts_forward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!link_tf || !idx || !cpts || !cnrm) return gqp_set_error(GQP_ERR_ARG, "gqp_contacts_forward: null pointer");
    int n = Bt * n_c;
    GQP_LAUNCH(contacts_forward_kernel, (n + 255) / 256, 256, 0, (cudaStream_t)stream, h->d_cand, h->d_cand_normal,
               h->d_cand_link, h->host.n_cand, h->host.n_links, link_tf, idx, n, n_c, cpts, cnrm);
    GQP_CHECK_LAUNCH("gqp_contacts_forward");
    return GQP_OK;
}

extern "C" int gqp_contacts_backward(const GqpHand *h, const int32_t *idx, const float *grad_cpts,
                                     const float *grad_cnrm, int Bt, int n_c, float *grad_link_tf, void *stream) {
    if (!h || Bt < 0 || n_c <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_contacts_backward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!idx || !grad_link_tf) return gqp_set_error(GQP_ERR_ARG, "gqp_contacts_backward: null pointer");
    GQP_LAUNCH(contacts_backward_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, h->d_cand, h->d_cand_normal,
               h->d_cand_link, h->host.n_cand, h->host.n_links, idx, grad_cpts, grad_cnrm, Bt, n_c, grad_link_tf);
    GQP_CHECK_LAUNCH("gqp_contacts_backward");
    return GQP_OK;
}

// ---- fused, warp-per-candidate entry points (what the engine launches) ----------------------------------------------
extern "C" int gqp_fk_contacts_forward(const GqpHand *h, const float *pose, const int32_t *idx, int Bt, int n_c,
                                       float *link_tf, float *cpts, float *cnrm, void *stream) {
    if (!h || Bt < 0 || n_c < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_contacts_forward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !link_tf || (idx && (!cpts || !cnrm))) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_contacts_forward: null pointer");
    if (h->host.n_frames > 32 || h->host.max_children > GQP_MAX_CHILDREN) {  // wide trees: the thread-per-candidate kernels
        int rc = gqp_fk_forward(h, pose, Bt, link_tf, stream);
        if (rc == GQP_OK && idx) rc = gqp_contacts_forward(h, link_tf, idx, Bt, n_c, cpts, cnrm, stream);
        return rc;
    }
    GQP_LAUNCH(fk_contacts_forward_warp_kernel, (Bt + FKW_WARPS - 1) / FKW_WARPS, FKW_WARPS * 32, 0, (cudaStream_t)stream,
               h->d_dev, h->d_cand, h->d_cand_normal, h->d_cand_link, pose, idx, Bt, n_c, link_tf, cpts, cnrm);
    GQP_CHECK_LAUNCH("gqp_fk_contacts_forward");
    return GQP_OK;
}

extern "C" int gqp_fk_contacts_backward(const GqpHand *h, const float *pose, const int32_t *idx, const float *grad_cpts,
                                        const float *grad_cnrm, float *grad_link_tf, const float *grad_pose_extra,
                                        int Bt, int n_c, float *grad_pose, void *stream) {
    if (!h || Bt < 0 || n_c < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_contacts_backward: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!pose || !grad_pose) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_contacts_backward: null pointer");
    if (h->host.n_frames > 32 || h->host.max_children > GQP_MAX_CHILDREN) {
        if (!grad_link_tf) return gqp_set_error(GQP_ERR_ARG, "gqp_fk_contacts_backward: grad_link_tf needed for this hand");
        int rc = idx ? gqp_contacts_backward(h, idx, grad_cpts, grad_cnrm, Bt, n_c, grad_link_tf, stream) : GQP_OK;
        if (rc == GQP_OK) rc = gqp_fk_backward(h, pose, grad_link_tf, grad_pose_extra, Bt, grad_pose, stream);
        return rc;
    }
    GQP_LAUNCH(fk_contacts_backward_warp_kernel, (Bt + FKW_WARPS - 1) / FKW_WARPS, FKW_WARPS * 32, 0, (cudaStream_t)stream,
               h->d_dev, h->d_cand, h->d_cand_normal, h->d_cand_link, pose, idx, grad_cpts, grad_cnrm, grad_link_tf,
               grad_pose_extra, Bt, n_c, grad_pose);
    GQP_CHECK_LAUNCH("gqp_fk_contacts_backward");
    return GQP_OK;
}
