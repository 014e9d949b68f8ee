// Exact nearest-face search over a BVH: returns the same winner as the brute-force sweep (lexicographic minimum of
// (dist2, original face index)).  A node is skipped only when the fp32 lower bound of its (inflated) box, shrunk by a
// relative 1e-5, still exceeds the current best — see gqp_bvh.cpp for why that can never hide a closer or tying face.
#pragma once
#include "gqp_hand.h"
#include "gqp_tri.cuh"

#define GQP_BVH_STACK 32

__device__ __forceinline__ float bvh_box_lb(const float4 &lo, const float4 &hi, V3 p) {
    float dx = fmaxf(0.f, fmaxf(lo.x - p.x, p.x - hi.x));
    float dy = fmaxf(0.f, fmaxf(lo.y - p.y, p.y - hi.y));
    float dz = fmaxf(0.f, fmaxf(lo.z - p.z, p.z - hi.z));
    return (dx * dx + dy * dy + dz * dz) * (1.0f - 1e-5f);
}

// nodes: GqpBvhNode array (shared or global), tris: (F,9) in leaf order, face_id: slot -> original index
template <bool EXACT>
__device__ __forceinline__ void bvh_query(const GqpBvhNode *nodes, const float *__restrict__ tris,
                                          const int32_t *__restrict__ face_id, V3 p, SdfBest &best) {
    int stack[GQP_BVH_STACK];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        const int ni = stack[--sp];
        const float4 *n4 = reinterpret_cast<const float4 *>(nodes + ni);
        const float4 lo = n4[0], hi = n4[1];
        if (bvh_box_lb(lo, hi, p) > best.d2) continue;
        const int a = __float_as_int(lo.w), b = __float_as_int(hi.w);
        if (b < 0) {
            for (int s = a; s < a - b; ++s) {
                const float *t = tris + 9 * (size_t)s;
                TriHit h = tri_closest<EXACT>(p, ld3(t), ld3(t + 3), ld3(t + 6));
                sdf_best_update(best, h, face_id[s]);
            }
        } else {
            const float4 *l4 = reinterpret_cast<const float4 *>(nodes + a);
            const float4 *r4 = reinterpret_cast<const float4 *>(nodes + b);
            float dl = bvh_box_lb(l4[0], l4[1], p), dr = bvh_box_lb(r4[0], r4[1], p);
            // push the farther child first so the nearer one is visited next
            if (dl <= dr) {
                if (dr <= best.d2) stack[sp++] = b;
                if (dl <= best.d2) stack[sp++] = a;
            } else {
                if (dl <= best.d2) stack[sp++] = a;
                if (dr <= best.d2) stack[sp++] = b;
            }
        }
    }
}
