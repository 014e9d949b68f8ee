// Point -> triangle closest point, the arithmetic of TorchSDF's forward kernel
// (thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu:131-165 helpers, :200-237 per-face body).
//
// Two instantiations:
//   EXACT = true : every product / sum is a separately rounded IEEE op (__fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn, which
//                  ptxas never contracts) in the reference's left-to-right order, rsqrt replaced by 1/sqrt.  Bit-identical
//                  to the reference kernel body compiled for the host with -ffp-contract=off (oracle/_ref).
//   EXACT = false: plain operators; nvcc contracts them into FFMA.  Same algorithm, ulp-level differences.  Used by the
//                  fused penetration kernel where throughput matters.
#pragma once
#include "gqp_common.cuh"

template <bool EXACT>
struct TriOps {
    static __device__ __forceinline__ float mul(float a, float b) { return EXACT ? __fmul_rn(a, b) : a * b; }
    static __device__ __forceinline__ float add(float a, float b) { return EXACT ? __fadd_rn(a, b) : a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return EXACT ? __fsub_rn(a, b) : a - b; }
    static __device__ __forceinline__ float div(float a, float b) { return EXACT ? __fdiv_rn(a, b) : a / b; }
    static __device__ __forceinline__ V3 sub(V3 a, V3 b) { return V3{sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)}; }
    static __device__ __forceinline__ V3 add(V3 a, V3 b) { return V3{add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)}; }
    static __device__ __forceinline__ V3 scl(V3 a, float s) { return V3{mul(a.x, s), mul(a.y, s), mul(a.z, s)}; }
    static __device__ __forceinline__ float dot(V3 a, V3 b) {
        return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z));
    }
    static __device__ __forceinline__ V3 cross(V3 a, V3 b) {
        return V3{sub(mul(a.y, b.z), mul(a.z, b.y)), sub(mul(a.z, b.x), mul(a.x, b.z)),
                  sub(mul(a.x, b.y), mul(a.y, b.x))};
    }
    static __device__ __forceinline__ float inv_sqrt(float a) {
        return EXACT ? __fdiv_rn(1.0f, __fsqrt_rn(a)) : rsqrtf(a);
    }
    // .cu:131-137
    static __device__ __forceinline__ float project_edge(V3 vertex, V3 edge, V3 p) {
        V3 pv = sub(p, vertex);
        float len = dot(edge, edge);
        return div(dot(pv, edge), len);
    }
    // .cu:160-165
    static __device__ __forceinline__ bool not_above(V3 vertex, V3 edge, V3 normal, V3 p) {
        return dot(cross(normal, edge), sub(p, vertex)) <= 0.f;
    }
};

__device__ __forceinline__ bool gqp_in_range(float a) { return a <= 1.f && a >= 0.f; }

struct TriHit {
    float d2;   // squared distance
    V3 c;       // closest point on the triangle
    V3 normal;  // un-normalised face normal cross(v1-v2, v1-v3)
};

template <bool EXACT>
__device__ __forceinline__ TriHit tri_closest(V3 p, V3 v1, V3 v2, V3 v3_) {
    typedef TriOps<EXACT> O;
    V3 e12 = O::sub(v2, v1), e23 = O::sub(v3_, v2), e31 = O::sub(v1, v3_);
    V3 normal = O::cross(O::sub(v1, v2), e31);
    float uab = O::project_edge(v1, e12, p);
    float uca = O::project_edge(v3_, e31, p);
    V3 c;
    if (uca > 1.f && uab < 0.f) {
        c = v1;
    } else {
        float ubc = O::project_edge(v2, e23, p);
        if (uab > 1.f && ubc < 0.f) {
            c = v2;
        } else if (ubc > 1.f && uca < 0.f) {
            c = v3_;
        } else if (gqp_in_range(uab) && O::not_above(v1, e12, normal, p)) {
            c = O::add(v1, O::scl(e12, uab));
        } else if (gqp_in_range(ubc) && O::not_above(v2, e23, normal, p)) {
            c = O::add(v2, O::scl(e23, ubc));
        } else if (gqp_in_range(uca) && O::not_above(v3_, e31, normal, p)) {
            c = O::add(v3_, O::scl(e31, uca));
        } else {  // project_plane .cu:139-148
            float inv_len = O::inv_sqrt(O::dot(normal, normal));
            V3 un = O::scl(normal, inv_len);
            float d = O::add(O::add(O::mul(O::sub(p.x, v1.x), un.x), O::mul(O::sub(p.y, v1.y), un.y)),
                             O::mul(O::sub(p.z, v1.z), un.z));
            c = O::sub(p, O::scl(un, d));
        }
    }
    TriHit h;
    V3 dv = O::sub(p, c);
    h.d2 = O::dot(dv, dv);
    h.c = c;
    h.normal = normal;
    return h;
}

// Running best of a first-minimum search, order independent: lexicographic (d2, face).
struct SdfBest {
    float d2;
    int face;
    V3 c, normal;
};
__device__ __forceinline__ void sdf_best_init(SdfBest &b) {
    b.d2 = __int_as_float(0x7f800000);
    b.face = 0x7fffffff;
    b.c = V3{0, 0, 0};
    b.normal = V3{0, 0, 0};
}
__device__ __forceinline__ void sdf_best_update(SdfBest &b, const TriHit &h, int face) {
    if (h.d2 < b.d2 || (h.d2 == b.d2 && face < b.face) || b.face == 0x7fffffff) {
        b.d2 = h.d2;
        b.face = face;
        b.c = h.c;
        b.normal = h.normal;
    }
}
// sign and unit normal of the winner (.cu:234-236)
template <bool EXACT>
__device__ __forceinline__ void sdf_finish(V3 p, const SdfBest &b, int &sign, V3 &unit) {
    typedef TriOps<EXACT> O;
    V3 dv = O::sub(p, b.c);
    sign = O::dot(dv, b.normal) >= 0.f ? 1 : -1;
    float inv = O::inv_sqrt(O::add(1e-16f, O::dot(dv, dv)));
    unit = O::scl(dv, inv);
}
