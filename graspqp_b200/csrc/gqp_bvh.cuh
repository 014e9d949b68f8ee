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

// Cheap lower bound of the squared distance point -> face from the face's pre-test record (gqp_bvh.cpp): with h the
// distance to the face's plane and t the in-plane distance to the circle that holds the face, d^2 >= h^2 + t^2.  An
// axis-aligned box around a tilted surface patch is as thick as the patch is wide, so box bounds alone let tens of
// faces through to the (expensive, exactly rounded) face test; this bound is tight to second order.
// Every rounding is resolved downwards (fp32 error of |v|^2 - h^2 and of h is below 1e-6 |v|^2), so the value never
// exceeds the true squared distance; bvh_face_skip() adds the slack that covers the rounding of the face test itself.
__device__ __forceinline__ float tri_lower_bound(const float4 q0, const float4 q1, V3 p, float &l1) {
    const float vx = p.x - q1.x, vy = p.y - q1.y, vz = p.z - q1.z;
    const float vv = vx * vx + vy * vy + vz * vz;
    const float h = vx * q0.x + vy * q0.y + vz * q0.z;
    const float ip2 = fmaxf(0.f, vv - h * h - 2e-6f * vv);
    const float t = fmaxf(0.f, sqrtf(ip2) - q0.w);
    l1 = fabsf(vx) + fabsf(vy) + fabsf(vz);
    return fmaxf(0.f, h * h - 4e-6f * vv) + t * t;
}
// true when the face cannot win or tie: its fp32 face-test result would exceed `best_d2` even after the worst-case
// rounding of that test (closest point off by a few ulp of the coordinate scale `scale`, i.e. d^2 off by
// <= 2 d * 1e-6 scale; l1 >= d)
__device__ __forceinline__ bool bvh_face_skip(float lb, float l1, float scale, float best_d2) {
    return lb * (1.0f - 1e-5f) - 8e-6f * scale * l1 > best_d2;
}

// nodes: GqpBvhNode array (shared or global), tris: (F,9) in leaf order, face_id: slot -> original index,
// pre: (F,8) pre-test records in leaf order, scale: max |coordinate| of mesh and query
template <bool EXACT>
__device__ __forceinline__ void bvh_query(const GqpBvhNode *nodes, const float *__restrict__ tris,
                                          const int32_t *__restrict__ face_id, const float *__restrict__ pre,
                                          float scale, V3 p, SdfBest &best) {
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
                const float4 *q = reinterpret_cast<const float4 *>(pre + 8 * (size_t)s);
                float l1;
                const float lbf = tri_lower_bound(q[0], q[1], p, l1);
                if (bvh_face_skip(lbf, l1, scale, best.d2)) continue;
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

// ---- warp-streamed form --------------------------------------------------------------------------------------------
// The same exact search, organised for SIMT efficiency when the lanes of a warp hold UNRELATED query points (contact
// points of different fingers / candidates): a warp owns a contiguous range of queries and its lanes pull the next
// query as soon as they finish one; every round of the outer loop is "descend inner nodes until a leaf is reached"
// (cheap, divergent) followed by "test that leaf's faces" (expensive, executed by all lanes that reached a leaf
// together).  The set of faces tested per query is exactly bvh_query's, so winners are identical.
// fetch(i, p, tree) loads query i and selects its tree; emit(i, p, best) stores the result.
struct BvhTree {
    const GqpBvhNode *nodes;
    const float *tris;
    const int32_t *face_id;
    const float *pre;
    float extent;
};
template <bool EXACT, class Fetch, class Emit>
__device__ __forceinline__ void bvh_query_stream(int q_begin, int q_end, int max_steps, Fetch fetch, Emit emit) {
    const int lane = threadIdx.x & 31;
    int stack[GQP_BVH_STACK];
    int sp = 0, node = -1, qi = -1;
    V3 p = V3{0.f, 0.f, 0.f};
    BvhTree tr = {nullptr, nullptr, nullptr, nullptr, 0.f};
    float scale = 0.f;
    SdfBest best;
    sdf_best_init(best);
    int next = q_begin;
    while (true) {
        // lanes without a query take the next ones of the warp's range, in lane order
        const bool need = qi < 0;
        const unsigned mask = __ballot_sync(GQP_FULL, need);
        if (need) {
            const int cand = next + __popc(mask & ((1u << lane) - 1u));
            if (cand < q_end) {
                qi = cand;
                fetch(qi, p, tr);
                scale = fmaxf(tr.extent, fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z))));
                sdf_best_init(best);
                sp = 0;
                node = 0;
            }
        }
        next += __popc(mask);
        if (!__any_sync(GQP_FULL, qi >= 0)) break;
        // phase 1: walk inner nodes until a leaf (or the end of the search).  `node` is the next node to examine, the
        // stack holds the others; a child entered directly keeps its box in registers (no dependent reload)
        int leaf_first = 0, leaf_count = 0;
        float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
        bool have_box = false;
        for (int steps = 0; node >= 0 && steps < max_steps; ++steps) {
            if (!have_box) {
                const float4 *n4 = reinterpret_cast<const float4 *>(tr.nodes + node);
                lo = n4[0];
                hi = n4[1];
                if (bvh_box_lb(lo, hi, p) > best.d2) {
                    node = sp > 0 ? stack[--sp] : -1;
                    continue;
                }
            }
            have_box = false;
            const int a = __float_as_int(lo.w), b = __float_as_int(hi.w);
            if (b < 0) {
                leaf_first = a;
                leaf_count = -b;
                node = sp > 0 ? stack[--sp] : -1;
                break;
            }
            const float4 *l4 = reinterpret_cast<const float4 *>(tr.nodes + a);
            const float4 *r4 = reinterpret_cast<const float4 *>(tr.nodes + b);
            const float4 llo = l4[0], lhi = l4[1], rlo = r4[0], rhi = r4[1];
            const float dl = bvh_box_lb(llo, lhi, p), dr = bvh_box_lb(rlo, rhi, p);
            const bool l_first = dl <= dr;  // the nearer child is examined next, the farther one after it
            if ((l_first ? dr : dl) <= best.d2) stack[sp++] = l_first ? b : a;
            if ((l_first ? dl : dr) <= best.d2) {
                node = l_first ? a : b;
                lo = l_first ? llo : rlo;
                hi = l_first ? lhi : rhi;
                have_box = true;
            } else {
                node = sp > 0 ? stack[--sp] : -1;
            }
        }
        // phase 2: the faces of the leaf, for all lanes that reached one: the cheap bounds of all its faces first
        // (coherent), then the exactly rounded face test of the few that can still win, nearest bound first
        unsigned todo = 0;
        for (int k = 0; k < leaf_count; ++k) {
            const float4 *q = reinterpret_cast<const float4 *>(tr.pre + 8 * (size_t)(leaf_first + k));
            float l1;
            const float lbf = tri_lower_bound(q[0], q[1], p, l1);
            if (!bvh_face_skip(lbf, l1, scale, best.d2)) todo |= 1u << k;
        }
        while (todo) {
            const int k = __ffs((int)todo) - 1;
            todo &= todo - 1;
            const int s = leaf_first + k;
            if (best.face != 0x7fffffff) {  // the bound may have tightened since the first pass
                const float4 *q = reinterpret_cast<const float4 *>(tr.pre + 8 * (size_t)s);
                float l1;
                const float lbf = tri_lower_bound(q[0], q[1], p, l1);
                if (bvh_face_skip(lbf, l1, scale, best.d2)) continue;
            }
            const float *t = tr.tris + 9 * (size_t)s;
            TriHit h = tri_closest<EXACT>(p, ld3(t), ld3(t + 3), ld3(t + 6));
            sdf_best_update(best, h, tr.face_id[s]);
        }
        if (qi >= 0 && node < 0) {  // nothing pending: the search of this query is complete
            emit(qi, p, best);
            qi = -1;
        }
    }
}
