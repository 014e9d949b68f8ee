// K1 — the TorchSDF contract: brute-force point -> triangle-mesh query, bit-identical to the reference arithmetic.
// Replaces thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu:176-270.
//
// Design (B200): a thread owns one point for the whole face sweep (the reference round-trips its running minimum
// through global memory per 512-face tile, .cu:245-250); faces stream through two shared-memory tiles filled by the TMA
// bulk-copy engine (one elected thread, mbarrier complete_tx) while the other tile is being consumed; every lane reads
// the same face (shared-memory broadcast).  Algorithmic HBM traffic: 12 B in + 36 B out per point, faces read once per
// block from L2.  Bound: FP32 issue (about 110 non-contracted ops per point-face pair).
#include "gqp_stage.cuh"
#include "gqp_tri.cuh"
#include "graspqp_b200.h"

#define SDF_TILE_F 512
#define SDF_THREADS 256

template <bool EXACT>
__global__ void __launch_bounds__(SDF_THREADS)
sdf_bruteforce_kernel(const float *__restrict__ points, const float *__restrict__ faces, int P, int F,
                      float *__restrict__ dist2, int32_t *__restrict__ sign, float *__restrict__ normals,
                      float *__restrict__ clst, int32_t *__restrict__ face_out) {
    __shared__ __align__(128) float tile[2][SDF_TILE_F * 9];
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x;
    const int n_tiles = (F + SDF_TILE_F - 1) / SDF_TILE_F;
    const bool aligned = (((uintptr_t)faces) & 15) == 0;
    if (tid == 0) {
        mbar_init(&bar[0]);
        mbar_init(&bar[1]);
    }
    __syncthreads();

    const int pi = blockIdx.x * SDF_THREADS + tid;
    V3 p = V3{0, 0, 0};
    if (pi < P) p = ld3(points + 3 * (size_t)pi);
    SdfBest best;
    sdf_best_init(best);
    uint32_t n_done[2] = {0, 0};

    // stage tile t into buffer b: bulk copy for the 16-byte multiple, plain loads for the remainder
    auto stage = [&](int t, int b) -> bool {
        int nf = min(SDF_TILE_F, F - t * SDF_TILE_F);
        uint32_t bytes = (uint32_t)nf * 36u;
        uint32_t bulk = aligned ? (bytes & ~15u) : 0u;
        const float *src = faces + (size_t)t * SDF_TILE_F * 9;
        if (bulk && tid == 0) bulk_load(&tile[b][0], src, bulk, &bar[b]);
        for (uint32_t i = bulk / 4 + tid; i < bytes / 4; i += SDF_THREADS) tile[b][i] = src[i];
        return bulk != 0;
    };

    bool pending = stage(0, 0);
    for (int t = 0; t < n_tiles; ++t) {
        const int b = t & 1;
        bool next_pending = false;
        if (t + 1 < n_tiles) next_pending = stage(t + 1, b ^ 1);
        if (pending) {
            mbar_wait(&bar[b], n_done[b]);
            n_done[b]++;
        }
        __syncthreads();
        const int nf = min(SDF_TILE_F, F - t * SDF_TILE_F);
        const float *tf = &tile[b][0];
        if (pi < P) {
#pragma unroll 2
            for (int f = 0; f < nf; ++f) {
                V3 v1 = ld3(tf + 9 * f), v2 = ld3(tf + 9 * f + 3), v3_ = ld3(tf + 9 * f + 6);
                TriHit h = tri_closest<EXACT>(p, v1, v2, v3_);
                sdf_best_update(best, h, t * SDF_TILE_F + f);
            }
        }
        __syncthreads();
        pending = next_pending;
    }
    if (pi < P) {
        int sg;
        V3 un;
        sdf_finish<EXACT>(p, best, sg, un);
        dist2[pi] = best.d2;
        sign[pi] = sg;
        st3(normals + 3 * (size_t)pi, un);
        st3(clst + 3 * (size_t)pi, best.c);
        if (face_out) face_out[pi] = best.face;
    }
}

// .cu:256-270: grad_points = 2 * grad_dist * (p - closest)
__global__ void sdf_backward_kernel(const float *__restrict__ grad_dist, const float *__restrict__ points,
                                    const float *__restrict__ clst, int P, float *__restrict__ grad_points) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 3 * P) return;
    float g = 2.f * grad_dist[i / 3];
    grad_points[i] = (points[i] - clst[i]) * g;
}

static void launch_sdf_exact(const float *points, const float *faces, int P, int F, float *dist2, int32_t *sign,
                             float *normals, float *clst, int32_t *face_out, cudaStream_t st) {
    int grid = (P + SDF_THREADS - 1) / SDF_THREADS;
    GQP_LAUNCH(sdf_bruteforce_kernel<true>, grid, SDF_THREADS, 0, st, points, faces, P, F, dist2, sign, normals, clst,
               face_out);
}

extern "C" int gqp_sdf_forward(const float *points, const float *face_verts, int P, int F, float *dist2,
                               int32_t *sign, float *normals, float *clst, void *stream) {
    if (P < 0 || F <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_sdf_forward: need P >= 0 and F > 0 (P=%d F=%d)", P, F);
    if (P == 0) return GQP_OK;
    if (!points || !face_verts || !dist2 || !sign || !normals || !clst)
        return gqp_set_error(GQP_ERR_ARG, "gqp_sdf_forward: null pointer");
    launch_sdf_exact(points, face_verts, P, F, dist2, sign, normals, clst, nullptr, (cudaStream_t)stream);
    GQP_CHECK_LAUNCH("gqp_sdf_forward");
    return GQP_OK;
}

extern "C" int gqp_sdf_backward(const float *grad_dist, const float *points, const float *clst, int P,
                                float *grad_points, void *stream) {
    if (P < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_sdf_backward: P < 0");
    if (P == 0) return GQP_OK;
    if (!grad_dist || !points || !clst || !grad_points) return gqp_set_error(GQP_ERR_ARG, "gqp_sdf_backward: null pointer");
    int n = 3 * P;
    GQP_LAUNCH(sdf_backward_kernel, (n + 255) / 256, 256, 0, (cudaStream_t)stream, grad_dist, points, clst, P,
               grad_points);
    GQP_CHECK_LAUNCH("gqp_sdf_backward");
    return GQP_OK;
}
