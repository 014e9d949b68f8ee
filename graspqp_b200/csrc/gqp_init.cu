// (f1) Initialisation / outlier re-initialisation ON THE DEVICE — graspqp/src/graspqp/core/initializations.py:15-193
// (`initialize_convex_hull`), which scripts/fit.py:421 runs on the HOST (trimesh + pytorch3d + Python loops) for all
// objects at every reset.  With the iteration at ~1 ms that host pass would dominate a 7000-iteration fit.
//
//   init_pool_kernel  a thread per pool point: area-weighted face of the object's convex hull (binary search of the
//                     cumulative areas), uniform barycentric point, pushed 1 cm along the face normal (:54-56).
//   init_fps_kernel   a CTA per object: farthest point sampling of batch_each of the pool_factor*batch_each pool points,
//                     start index 0, first maximum (pytorch3d.ops.sample_farthest_points, :60-62).  The running minimum
//                     distances live in global memory (L2 resident: 16 + 4 bytes per pool point), one block-wide
//                     arg-max per pick.
//   init_pose_kernel  a thread per candidate: inward normal, look_at (:84-120), Euler jitter (:122-145), translation
//                     (:147-150), truncated-normal joints (:164-176), contact indices (:190-192).
// Noise: injected arrays of uniforms (parity runs) or Philox4x32-10 keyed (seed; counter = reset number).
// Arithmetic of the pool points and of the FPS distances is not FMA-contracted, in the order of oracle/init.py, so the
// picked indices can be compared bit for bit.
#include "gqp_hand.h"
#include "gqp_philox.cuh"
#include "graspqp_b200.h"

#ifdef GQP_EMU
#define FPS_THREADS 64
#else
#define FPS_THREADS 1024
#endif

struct GqpInit {
    int n_obj = 0;
    std::vector<int32_t> off_host;  // (n_obj+1) first hull face of every object
    int32_t *d_off = nullptr;
    float *d_tris = nullptr;  // (H,9)
    float *d_nrm = nullptr;   // (H,3) unit outward normals
    float *d_cdf = nullptr;   // (H) cumulative area fraction within the object
};

struct InitDev {  // by-value kernel argument
    float distance[2], rotate[2], pitch[2], tilt[2];  // lower, upper - lower
    float inflate;
    float fwd[3], up[3];
    int32_t dof;
    // truncated normal per joint (torch.nn.init.trunc_normal_): x = clamp(erfinv(u * tn_w + tn_lo) * tn_s + tn_mu, a, b)
    float tn_lo[GQP_MAX_DOFS], tn_w[GQP_MAX_DOFS], tn_s[GQP_MAX_DOFS], tn_mu[GQP_MAX_DOFS], tn_a[GQP_MAX_DOFS],
        tn_b[GQP_MAX_DOFS];
};

// inverse error function, single precision (M. Giles, "Approximating the erfinv function", 2010): relative error ~4e-7
__device__ __forceinline__ float gqp_erfinv(float x) {
    float w = -logf((1.0f - x) * (1.0f + x)), p;
    if (w < 5.0f) {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = 3.43273939e-07f + p * w; p = -3.5233877e-06f + p * w; p = -4.39150654e-06f + p * w;
        p = 0.00021858087f + p * w; p = -0.00125372503f + p * w; p = -0.00417768164f + p * w;
        p = 0.246640727f + p * w; p = 1.50140941f + p * w;
    } else {
        w = sqrtf(w) - 3.0f;
        p = -0.000200214257f;
        p = 0.000100950558f + p * w; p = 0.00134934322f + p * w; p = -0.00367342844f + p * w;
        p = 0.00573950773f + p * w; p = -0.0076224613f + p * w; p = 0.00943887047f + p * w;
        p = 1.00167406f + p * w; p = 2.83297682f + p * w;
    }
    return p * x;
}

__global__ void __launch_bounds__(256)
init_pool_kernel(const int32_t *__restrict__ off, const float *__restrict__ tris, const float *__restrict__ nrm,
                 const float *__restrict__ cdf, int n_obj, int M, float inflate, const float *__restrict__ u_pool,
                 uint64_t seed, uint64_t counter, float4 *__restrict__ pool, float *__restrict__ mind) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_obj * M) return;
    const int o = (int)(i / M);
    float u0, u1, u2;
    if (u_pool) {
        u0 = u_pool[3 * i]; u1 = u_pool[3 * i + 1]; u2 = u_pool[3 * i + 2];
    } else {
        Philox r = philox4x32((uint32_t)i, (uint32_t)(i >> 32) | 0x10000000u, (uint32_t)counter, (uint32_t)(counter >> 32),
                              (uint32_t)seed, (uint32_t)(seed >> 32));
        u0 = u01(r.c[0]); u1 = u01(r.c[1]); u2 = u01(r.c[2]);
    }
    // first face whose cumulative area fraction exceeds u0 (numpy.searchsorted(side="right"), clamped)
    int lo = off[o], hi = off[o + 1];
    const int last = hi - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] > u0) hi = mid; else lo = mid + 1;
    }
    const int f = min(lo, last);
    const float s = __fsqrt_rn(u1);
    const float w0 = __fsub_rn(1.0f, s), w1 = __fmul_rn(s, __fsub_rn(1.0f, u2)), w2 = __fmul_rn(s, u2);
    const float *t = tris + 9 * (size_t)f;
    float p[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float q = __fadd_rn(__fadd_rn(__fmul_rn(w0, t[k]), __fmul_rn(w1, t[3 + k])), __fmul_rn(w2, t[6 + k]));
        p[k] = __fadd_rn(q, __fmul_rn(inflate, nrm[3 * (size_t)f + k]));
    }
    pool[i] = make_float4(p[0], p[1], p[2], __int_as_float(f));
    mind[i] = __int_as_float(0x7f800000);
}

// farthest point sampling, one CTA per object
__global__ void __launch_bounds__(FPS_THREADS)
init_fps_kernel(const float4 *__restrict__ pool, float *__restrict__ mind, int M, int K, int32_t *__restrict__ sel) {
    __shared__ float s_val[FPS_THREADS / 32];
    __shared__ int s_idx[FPS_THREADS / 32];
    __shared__ int s_cur;
    const int o = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 *P = pool + (size_t)o * M;
    float *D = mind + (size_t)o * M;
    int cur = 0;
    for (int k = 0; k < K; ++k) {
        if (threadIdx.x == 0) sel[(size_t)o * K + k] = cur;
        const float4 c = P[cur];
        float bv = -1.0f;
        int bi = 0x7fffffff;
        for (int j = threadIdx.x; j < M; j += FPS_THREADS) {
            const float4 q = P[j];
            const float dx = __fsub_rn(q.x, c.x), dy = __fsub_rn(q.y, c.y), dz = __fsub_rn(q.z, c.z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const float m = fminf(D[j], d2);
            D[j] = m;
            if (m > bv) { bv = m; bi = j; }  // ascending j per thread: the first maximum survives
        }
#pragma unroll
        for (int w = 16; w > 0; w >>= 1) {
            const float ov = __shfl_xor_sync(GQP_FULL, bv, w);
            const int oi = __shfl_xor_sync(GQP_FULL, bi, w);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            bv = lane < FPS_THREADS / 32 ? s_val[lane] : -1.0f;
            bi = lane < FPS_THREADS / 32 ? s_idx[lane] : 0x7fffffff;
#pragma unroll
            for (int w = 16; w > 0; w >>= 1) {
                const float ov = __shfl_xor_sync(GQP_FULL, bv, w);
                const int oi = __shfl_xor_sync(GQP_FULL, bi, w);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) s_cur = bi;
        }
        __syncthreads();
        cur = s_cur;
        __syncthreads();
    }
}

// a thread per candidate: the pose of initializations.py:79-193 from its picked pool point
__global__ void __launch_bounds__(128)
init_pose_kernel(InitDev P, const float4 *__restrict__ pool, const int32_t *__restrict__ sel,
                 const float *__restrict__ nrm, int n_obj, int batch_each, int M, int n_c, int n_cand,
                 const float *__restrict__ u_pose, const int32_t *__restrict__ idx_rand, uint64_t seed, uint64_t counter,
                 float *__restrict__ pose_init, int32_t *__restrict__ idx_init) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_obj * batch_each) return;
    const int o = b / batch_each, D = 9 + P.dof, NU = 4 + P.dof;
    const float4 q = pool[(size_t)o * M + sel[b]];
    const int f = __float_as_int(q.w);
    const V3 p = V3{q.x, q.y, q.z};
    const V3 n = V3{-nrm[3 * (size_t)f], -nrm[3 * (size_t)f + 1], -nrm[3 * (size_t)f + 2]};  // :63-79 inward normal
    auto uni = [&](int k) -> float {
        if (u_pose) return u_pose[(size_t)b * NU + k];
        Philox r = philox4x32((uint32_t)b, 0x20000000u | (uint32_t)(k >> 2), (uint32_t)counter, (uint32_t)(counter >> 32),
                              (uint32_t)seed, (uint32_t)(seed >> 32));
        return u01(r.c[k & 3]);
    };
    const float distance = P.distance[0] + P.distance[1] * uni(0), rotate = P.rotate[0] + P.rotate[1] * uni(1),
                pitch = P.pitch[0] + P.pitch[1] * uni(2), tilt = P.tilt[0] + P.tilt[1] * uni(3);
    // look_at(p, p + n, up_vector = up axis, forward_vector = forward axis)  (:84-120)
    const V3 tgt = p + n;
    V3 fw = p - tgt;
    fw = fw * (1.0f / sqrtf(dot3(fw, fw)));
    const V3 base_up = V3{P.up[0], P.up[1], P.up[2]}, fv = V3{P.fwd[0], P.fwd[1], P.fwd[2]};
    const V3 upv = fabsf(dot3(base_up, fw)) < 0.95f ? base_up : V3{0.f, 1.f, 0.f};
    V3 right = cross3(upv, fw);
    right = right * (1.0f / sqrtf(dot3(right, right)));
    const V3 up2 = cross3(fw, right);
    const V3 b1 = cross3(fv, base_up) * -1.0f;
    // orientation = [fw up2 right] (columns), basis = [fv b1 base_up] (columns); Rg = orientation @ basis
    float Rg[9];
    const float oc[3][3] = {{fw.x, up2.x, right.x}, {fw.y, up2.y, right.y}, {fw.z, up2.z, right.z}};
    const float bs[3][3] = {{fv.x, b1.x, base_up.x}, {fv.y, b1.y, base_up.y}, {fv.z, b1.z, base_up.z}};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Rg[3 * i + j] = oc[i][0] * bs[0][j] + oc[i][1] * bs[1][j] + oc[i][2] * bs[2][j];
    // rotation_local = euler2mat(tilt, pitch, rotate, 'rxyz') = Rx(tilt) Ry(pitch) Rz(rotate)  (:138-145)
    const float ci = cosf(tilt), si = sinf(tilt), cj = cosf(pitch), sj = sinf(pitch), ck = cosf(rotate), sk = sinf(rotate);
    const float Rl[9] = {cj * ck, -cj * sk, sj, si * sj * ck + ci * sk, -si * sj * sk + ci * ck, -si * cj,
                         -ci * sj * ck + si * sk, ci * sj * sk + si * ck, ci * cj};
    float R[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R[3 * i + j] = Rg[3 * i] * Rl[j] + Rg[3 * i + 1] * Rl[3 + j] + Rg[3 * i + 2] * Rl[6 + j];
    float *out = pose_init + (size_t)b * D;
    out[0] = p.x - distance * n.x; out[1] = p.y - distance * n.y; out[2] = p.z - distance * n.z;  // :147
    out[3] = R[0]; out[4] = R[3]; out[5] = R[6];  // rotation.transpose(1,2)[:, :2]: the first two COLUMNS of R  (:187)
    out[6] = R[1]; out[7] = R[4]; out[8] = R[7];
    for (int j = 0; j < P.dof; ++j) {  // :164-176
        const float x = gqp_erfinv(uni(4 + j) * P.tn_w[j] + P.tn_lo[j]) * P.tn_s[j] + P.tn_mu[j];
        out[9 + j] = fminf(fmaxf(x, P.tn_a[j]), P.tn_b[j]);
    }
    for (int c = 0; c < n_c; ++c) {  // :190-192
        int32_t r;
        if (idx_rand) r = idx_rand[(size_t)b * n_c + c];
        else {
            Philox rr = philox4x32((uint32_t)b, 0x30000000u | (uint32_t)c, (uint32_t)counter, (uint32_t)(counter >> 32),
                                   (uint32_t)seed, (uint32_t)(seed >> 32));
            r = (int32_t)(((uint64_t)rr.c[0] * (uint64_t)n_cand) >> 32);
        }
        idx_init[(size_t)b * n_c + c] = r;
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
#include <cmath>
extern "C" int gqp_init_create(const float *hull_tris_host, const int32_t *hull_offset_host, int n_obj, GqpInit **out) {
    if (!hull_tris_host || !hull_offset_host || n_obj <= 0 || !out) return gqp_set_error(GQP_ERR_ARG, "gqp_init_create: bad argument");
    GqpInit *I = new GqpInit();
    I->n_obj = n_obj;
    I->off_host.assign(hull_offset_host, hull_offset_host + n_obj + 1);
    const int H = I->off_host[n_obj];
    std::vector<float> nrm(3 * (size_t)H), cdf(H);
    for (int o = 0; o < n_obj; ++o) {
        const int f0 = I->off_host[o], f1 = I->off_host[o + 1];
        if (f1 - f0 < 4) { delete I; return gqp_set_error(GQP_ERR_ARG, "gqp_init_create: object %d has no hull", o); }
        std::vector<double> acc(f1 - f0);
        double total = 0.0;
        for (int f = f0; f < f1; ++f) {
            const float *t = hull_tris_host + 9 * (size_t)f;
            const double e1[3] = {(double)t[3] - t[0], (double)t[4] - t[1], (double)t[5] - t[2]};
            const double e2[3] = {(double)t[6] - t[0], (double)t[7] - t[1], (double)t[8] - t[2]};
            const double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            const double l = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            if (!(l > 0.0)) { delete I; return gqp_set_error(GQP_ERR_ARG, "gqp_init_create: degenerate hull face %d", f); }
            for (int k = 0; k < 3; ++k) nrm[3 * (size_t)f + k] = (float)(n[k] / l);
            total += 0.5 * l;
            acc[f - f0] = total;
        }
        for (int f = f0; f < f1; ++f) cdf[f] = (float)(acc[f - f0] / total);
    }
    I->d_off = (int32_t *)gqp_dev_alloc_upload(I->off_host.data(), sizeof(int32_t) * (n_obj + 1));
    I->d_tris = (float *)gqp_dev_alloc_upload(hull_tris_host, sizeof(float) * 9 * (size_t)H);
    I->d_nrm = (float *)gqp_dev_alloc_upload(nrm.data(), sizeof(float) * nrm.size());
    I->d_cdf = (float *)gqp_dev_alloc_upload(cdf.data(), sizeof(float) * cdf.size());
    if (!I->d_off || !I->d_tris || !I->d_nrm || !I->d_cdf) return gqp_set_error(GQP_ERR_CUDA, "gqp_init_create: allocation failed");
    *out = I;
    return GQP_OK;
}
extern "C" void gqp_init_destroy(GqpInit *I) {
    if (!I) return;
    gqp_dev_free(I->d_off); gqp_dev_free(I->d_tris); gqp_dev_free(I->d_nrm); gqp_dev_free(I->d_cdf);
    delete I;
}
extern "C" uint64_t gqp_init_scratch_bytes(int n_obj, int batch_each, int pool_factor) {
    if (n_obj <= 0 || batch_each <= 0 || pool_factor <= 0) return 0;
    const uint64_t M = (uint64_t)pool_factor * batch_each;
    return (uint64_t)n_obj * M * (16 + 4) + (uint64_t)n_obj * batch_each * 4 + 64;
}
extern "C" int gqp_init_convex_hull(const GqpInit *I, const GqpHand *h, const GqpInitParams *p, int batch_each, int n_c,
                                    const float *u_pool, const float *u_pose, const int32_t *idx_rand, uint64_t seed,
                                    uint64_t counter, void *scratch, float *pose_init, int32_t *idx_init,
                                    int32_t *picked_out, void *stream) {
    if (!I || !h || !p || batch_each <= 0 || n_c <= 0 || p->pool_factor <= 0)
        return gqp_set_error(GQP_ERR_ARG, "gqp_init_convex_hull: bad argument");
    if (!scratch || !pose_init || !idx_init) return gqp_set_error(GQP_ERR_ARG, "gqp_init_convex_hull: null pointer");
    if (((uintptr_t)scratch & 15) != 0) return gqp_set_error(GQP_ERR_ARG, "gqp_init_convex_hull: scratch must be 16-byte aligned");
    const int n_obj = I->n_obj, dof = h->host.n_dofs;
    const long long M = (long long)p->pool_factor * batch_each;
    if (M * n_obj > 0x7fffffffLL) return gqp_set_error(GQP_ERR_LIMIT, "gqp_init_convex_hull: pool too large");
    InitDev P{};
    auto span = [](float lo, float hi, float *o) { o[0] = lo; o[1] = hi - lo; };
    span(p->distance_lower, p->distance_upper, P.distance); span(p->rotate_lower, p->rotate_upper, P.rotate);
    span(p->pitch_lower, p->pitch_upper, P.pitch); span(p->tilt_lower, p->tilt_upper, P.tilt);
    P.inflate = p->inflate;
    for (int k = 0; k < 3; ++k) { P.fwd[k] = p->forward_axis[k]; P.up[k] = p->up_axis[k]; }
    P.dof = dof;
    for (int j = 0; j < dof; ++j) {  // torch.nn.init.trunc_normal_: bounds of the uniform draw in double, as torch does
        const double lo = h->host.joints_lower[j], hi = h->host.joints_upper[j];
        const double mu = std::fmin(std::fmax((double)p->default_state[j], lo), hi);  // :164-165 clamp_
        const double sd = (double)p->jitter_strength * (hi - lo), a = lo - 1e-6, b = hi + 1e-6;
        auto ncdf = [](double x) { return (1.0 + std::erf(x / std::sqrt(2.0))) / 2.0; };
        const double l = 2 * ncdf((a - mu) / sd) - 1, u = 2 * ncdf((b - mu) / sd) - 1;
        P.tn_lo[j] = (float)l; P.tn_w[j] = (float)(u - l); P.tn_s[j] = (float)(sd * std::sqrt(2.0));
        P.tn_mu[j] = (float)mu; P.tn_a[j] = (float)a; P.tn_b[j] = (float)b;
    }
    float4 *pool = (float4 *)scratch;
    float *mind = (float *)(pool + (size_t)n_obj * M);
    int32_t *sel = picked_out ? picked_out : (int32_t *)(mind + (size_t)n_obj * M);
    cudaStream_t st = (cudaStream_t)stream;
    GQP_LAUNCH(init_pool_kernel, (unsigned)((n_obj * M + 255) / 256), 256, 0, st, I->d_off, I->d_tris, I->d_nrm, I->d_cdf,
               n_obj, (int)M, P.inflate, u_pool, seed, counter, pool, mind);
    GQP_CHECK_LAUNCH("gqp_init_convex_hull/pool");
    GQP_LAUNCH(init_fps_kernel, n_obj, FPS_THREADS, 0, st, pool, mind, (int)M, batch_each, sel);
    GQP_CHECK_LAUNCH("gqp_init_convex_hull/fps");
    GQP_LAUNCH(init_pose_kernel, (n_obj * batch_each + 127) / 128, 128, 0, st, P, pool, sel, I->d_nrm, n_obj, batch_each,
               (int)M, n_c, h->host.n_cand, u_pose, idx_rand, seed, counter, pose_init, idx_init);
    GQP_CHECK_LAUNCH("gqp_init_convex_hull/pose");
    return GQP_OK;
}
