/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement of TorchSDF's brute-force point -> triangle-mesh query, the arithmetic of
 *   /root/reference/thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu:176-254 (forward)
 *   and :256-270 (backward).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * Pinned against: oracle/_ref/libtorchsdf_ref.so (the reference kernel body itself, host-compiled by
 * oracle/build_ref.sh) in tests/test_oracle_sdf.py, and against the self-consistency law of
 * thirdparty/TorchSDF/tests/normal.py:36-39.
 *
 * Semantics kept: per face the 7-region closest point (vertex regions first, then edges in order 12,23,31 with the
 * "not above" test, else plane projection); squared distance; sign = (dot(p-c, cross(v1-v2, v1-v3)) >= 0) ? +1 : -1;
 * normal = (p-c) * rsqrt(1e-16 + |p-c|^2); winner = FIRST minimum in face order (strict >).
 * The reference tiles faces by 512 and keeps the running best in global memory with a strict compare, which is the
 * same first-minimum rule, so no tiling is restated here.
 * Build with -ffp-contract=off: every product and sum is rounded separately (the GPU reference contracts to FMA where
 * nvcc chooses; differences are ulp-level and covered by the 1e-5 absolute tolerance of the north star).
 */
#include <math.h>
#include <stdint.h>

typedef struct { float x, y, z; } v3;

static inline v3 sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline v3 add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline v3 scl(v3 a, float s) { v3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 a, v3 b) {
    v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
    return r;
}
/* .cu:131-137 */
static inline float project_edge(v3 vertex, v3 edge, v3 p) { return dot(sub(p, vertex), edge) / dot(edge, edge); }
/* .cu:160-165 */
static inline int not_above(v3 vertex, v3 edge, v3 normal, v3 p) { return dot(cross(normal, edge), sub(p, vertex)) <= 0.f; }
static inline int in_range(float a) { return a <= 1.f && a >= 0.f; }

/* one (point, face) test: closest point on the triangle; returns squared distance. .cu:200-237 */
static inline float tri_closest(v3 p, v3 v1, v3 v2, v3 v3_, v3 *clst, v3 *nrm_out) {
    v3 e12 = sub(v2, v1), e23 = sub(v3_, v2), e31 = sub(v1, v3_);
    v3 normal = cross(sub(v1, v2), e31);
    float uab = project_edge(v1, e12, p);
    float uca = project_edge(v3_, e31, p);
    v3 c;
    if (uca > 1.f && uab < 0.f) {
        c = v1;
    } else {
        float ubc = project_edge(v2, e23, p);
        if (uab > 1.f && ubc < 0.f) c = v2;
        else if (ubc > 1.f && uca < 0.f) c = v3_;
        else if (in_range(uab) && not_above(v1, e12, normal, p)) c = add(v1, scl(e12, uab));
        else if (in_range(ubc) && not_above(v2, e23, normal, p)) c = add(v2, scl(e23, ubc));
        else if (in_range(uca) && not_above(v3_, e31, normal, p)) c = add(v3_, scl(e31, uca));
        else { /* project_plane .cu:139-148 */
            float inv_len = 1.0f / sqrtf(dot(normal, normal));
            v3 un = scl(normal, inv_len);
            float d = (p.x - v1.x) * un.x + (p.y - v1.y) * un.y + (p.z - v1.z) * un.z;
            c = sub(p, scl(un, d));
        }
    }
    *clst = c;
    *nrm_out = normal;
    v3 dv = sub(p, c);
    return dot(dv, dv);
}

/* points (P,3), faces (F,3,3); outputs dist2 (P), sign (P) int32, normals (P,3), clst (P,3), face (P) int32 (extra:
 * the winning face index, not part of the reference contract, used by tests of the culled kernels). */
void sdf_oracle_forward(const float *points, const float *faces, int P, int F, float *dist2, int32_t *sign,
                        float *normals, float *clst, int32_t *face_idx) {
    const v3 *pts = (const v3 *)points;
    const v3 *fv = (const v3 *)faces;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        v3 p = pts[i];
        float best = INFINITY;
        int bsign = 0, bface = -1;
        v3 bn = {0, 0, 0}, bc = {0, 0, 0};
        for (int f = 0; f < F; ++f) {
            v3 c, normal;
            float d = tri_closest(p, fv[3 * f], fv[3 * f + 1], fv[3 * f + 2], &c, &normal);
            if (f == 0 || best > d) {
                v3 dv = sub(p, c);
                best = d;
                bsign = dot(dv, normal) >= 0.f ? 1 : -1;
                bn = scl(dv, 1.0f / sqrtf(1e-16f + dot(dv, dv)));
                bc = c;
                bface = f;
            }
        }
        dist2[i] = best;
        sign[i] = bsign;
        ((v3 *)normals)[i] = bn;
        ((v3 *)clst)[i] = bc;
        if (face_idx) face_idx[i] = bface;
    }
}

/* .cu:256-270: grad_points = 2 * grad_dist * (p - closest) */
void sdf_oracle_backward(const float *grad_dist, const float *points, const float *clst, int P, float *grad_points) {
    for (int i = 0; i < P; ++i) {
        float g = 2.f * grad_dist[i];
        for (int k = 0; k < 3; ++k) grad_points[3 * i + k] = (points[3 * i + k] - clst[3 * i + k]) * g;
    }
}

/* bench.py's reference arm runs under torchrun, which exports OMP_NUM_THREADS=1: set the team size explicitly */
#ifdef _OPENMP
#include <omp.h>
void sdf_oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int sdf_oracle_get_threads(void) { return omp_get_max_threads(); }
#else
void sdf_oracle_set_threads(int n) { (void)n; }
int sdf_oracle_get_threads(void) { return 1; }
#endif
