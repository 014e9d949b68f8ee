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
// The same record describes a LEAF (a patch of <= 8 faces): q1.w is then the half-thickness of the slab around the
// patch's mean plane that holds all its vertices (0 for a single face), and d^2 >= max(0, |h| - q1.w)^2 + t^2.
// Every rounding is resolved downwards (fp32 error of |v|^2 - h^2 and of h is below 1e-6 |v|^2), so the value never
// exceeds the true squared distance; bvh_face_skip() adds the slack that covers the rounding of the face test itself.
__device__ __forceinline__ float tri_lower_bound(const float4 q0, const float4 q1, V3 p, float &l1) {
    const float vx = p.x - q1.x, vy = p.y - q1.y, vz = p.z - q1.z;
    const float vv = vx * vx + vy * vy + vz * vz;
    const float h = vx * q0.x + vy * q0.y + vz * q0.z;
    const float ip2 = fmaxf(0.f, vv - h * h - 2e-6f * vv);
    const float t = fmaxf(0.f, sqrtf(ip2) - q0.w);
    const float hh = fmaxf(0.f, fabsf(h) - q1.w);
    l1 = fabsf(vx) + fabsf(vy) + fabsf(vz);
    return fmaxf(0.f, hh * hh - 4e-6f * vv) + t * t;
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

// ---- gang form: 8 lanes per query over the 8-wide tree ---------------------------------------------------------------
// The same exact search, organised for SIMT efficiency when the lanes of a warp hold UNRELATED query points (contact
// points of different fingers / candidates).  A thread-per-query walk of a binary tree leaves ~6 of 32 lanes busy (every
// lane is at a different node).  Here a gang of 8 lanes owns one query: the binary tree is collapsed on the host into
// nodes with up to 8 children (gqp_bvh_widen), every lane tests one child box of the node being expanded and one face
// of the leaf being examined, and the gang shares one traversal stack in shared memory (nearest child on top)
// and one pruning bound (the minimum of the lanes' best squared distances).  Each lane keeps the lexicographic best of
// the faces IT tested; the gang's winner is the lexicographic minimum of those, i.e. the brute-force sweep's winner.
// All warp collectives are executed by all 32 lanes (the 4 gangs of a warp loop in lockstep), as gqp_emu.h requires.
//
// Wide node w = 8 consecutive GqpBvhNode records (one per child): box, then
//   b < 0 : leaf child, a = first slot, -b = count (<= 8)      b == 1 : inner child, a = its wide-node index
//   b == 0: no child
#define GQP_GANG 8
#if defined(GQP_EMU) && defined(GQP_BVH_STATS)
#include <atomic>
inline std::atomic<long> g_bvh_stats[8];
extern "C" __attribute__((used)) inline void gqp_bvh_stats_read(long *out) { for (int i = 0; i < 8; ++i) { out[i] = g_bvh_stats[i]; g_bvh_stats[i] = 0; } }
#define BVH_STAT(i) do { if (gl == 0) g_bvh_stats[i]++; } while (0)
#define BVH_STAT_L(i) do { g_bvh_stats[i]++; } while (0)
#else
#define BVH_STAT(i)
#define BVH_STAT_L(i)
#endif
#define GQP_GANG_STACK 64  // >= 7 * depth + 1, checked at gqp_mesh_create
struct BvhTree {
    const GqpBvhNode *nodes;  // wide records
    const float *wpre;        // (8 per wide node, 8 floats) patch record of a leaf child (tri_lower_bound)
    const float *tris;
    const int32_t *face_id;
    const float *pre;
    float extent;
};
// A warp owns the queries [q_begin, q_end); a gang that finishes a query takes the warp's next one, so the 4 gangs of a
// warp stay busy although their searches differ in length.  fetch(i, p, tree) loads query i (executed by the 8 lanes
// of the gang).  done(i, p, tree, winner) is called on all 8 lanes when a search completes, with the winner broadcast;
// it stores the result and returns false, or replaces p / tree and returns true to run a follow-up search under the
// same query index (the hand-penetration search walks several link meshes per sample).
// stk: this gang's GQP_GANG_STACK entries (code, lower bound bits) in shared memory.
template <bool EXACT, class Fetch, class Done>
__device__ __forceinline__ void bvh_query_gang(int q_begin, int q_end, int2 *stk, Fetch fetch, Done done) {
    const int lane = threadIdx.x & 31, gl = lane & (GQP_GANG - 1), g0 = lane & ~(GQP_GANG - 1);
    const float INF = __int_as_float(0x7f800000);
    BvhTree tr = {nullptr, nullptr, nullptr, nullptr, nullptr, 0.f};
    V3 p = V3{0.f, 0.f, 0.f};
    float scale = 0.f, bound = INF;
    SdfBest best;
    sdf_best_init(best);
    int sp = 0, qi = -1, next = q_begin;
    while (true) {
        // idle gangs take the next queries of the warp's range, in gang order
        const bool need = qi < 0;
        const unsigned gm = __ballot_sync(GQP_FULL, need && gl == 0);
        if (need) {
            const int cand = next + __popc(gm & ((1u << g0) - 1u));
            if (cand < q_end) {
                qi = cand;
                fetch(qi, p, tr);
                scale = fmaxf(tr.extent, fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z))));
                sdf_best_init(best);
                bound = INF;
                if (gl == 0) stk[0] = make_int2(0, 0);  // root, lower bound 0
                sp = 1;
            }
        }
        next += __popc(gm);
        if (!__any_sync(GQP_FULL, qi >= 0)) break;
        __syncwarp();
        // phase 1: pop and expand wide nodes until the gang holds a leaf (or its stack is empty)
        int leaf_first = 0, leaf_count = 0;
        while (true) {
            const bool want = sp > 0 && leaf_count == 0;
            if (!__any_sync(GQP_FULL, want)) break;
            bool expand = false;
            int code = 0;
            if (want) {
                int2 e = stk[--sp];
                BVH_STAT(0);
                // entries pushed before the bound tightened: drop them here, a compare each, not a round of the warp
                while (__int_as_float(e.y) > bound && sp > 0) e = stk[--sp];
                if (__int_as_float(e.y) <= bound) {
                    if (e.x < 0) {
                        const int v = -(e.x + 1);
                        leaf_first = v >> 3;
                        leaf_count = (v & 7) + 1;
                    } else {
                        expand = true;
                        code = e.x;
                        BVH_STAT(1);
                    }
                }
            }
            __syncwarp();  // all lanes have read the popped entry before its slot is reused
            float d = INF;
            int ccode = 0;
            bool alive = false;
            if (expand) {
                const float4 *n4 = reinterpret_cast<const float4 *>(tr.nodes + GQP_GANG * (size_t)code + gl);
                const float4 lo = n4[0], hi = n4[1];
                const int a = __float_as_int(lo.w), b = __float_as_int(hi.w);
                if (b != 0) {
                    d = bvh_box_lb(lo, hi, p);
                    alive = d <= bound;
                    ccode = b < 0 ? -(a * 8 + (-b - 1)) - 1 : a;
                    if (alive && b < 0) {
                        // a leaf is a nearly flat patch: the slab-and-circle bound is far tighter than its box.  No face
                        // of the patch is farther from p than l1 + radius, which scales the face test's rounding slack
                        const float4 *q = reinterpret_cast<const float4 *>(tr.wpre + 8 * (GQP_GANG * (size_t)code + gl));
                        const float4 q0 = q[0], q1 = q[1];
                        float l1;
                        const float lbf = tri_lower_bound(q0, q1, p, l1);
                        const float dl = lbf * (1.0f - 1e-5f) - 8e-6f * scale * (l1 + q0.w);
                        d = fmaxf(d, dl);
                        alive = d <= bound;
                    }
                }
            }
            // survivors go on the stack with the nearest one on top (examined next); the others keep lane order — they are
            // re-checked against the bound when popped.  Ordering key: the bound's bits (d >= 0) with the lane in the low 3
            const unsigned am = (__ballot_sync(GQP_FULL, alive) >> g0) & 0xffu;
            int kmin = alive ? ((__float_as_int(d) & ~7) | gl) : 0x7fffffff;
#pragma unroll
            for (int o = 1; o < GQP_GANG; o <<= 1) kmin = min(kmin, __shfl_xor_sync(GQP_FULL, kmin, o));
            const int nearest = kmin & 7;
            if (alive) {
                const unsigned below = am & ~(1u << nearest) & ((1u << gl) - 1u);
                stk[gl == nearest ? sp + __popc(am) - 1 : sp + __popc(below)] = make_int2(ccode, __float_as_int(d));
            }
            sp += __popc(am);
            if (gl == 0) { BVH_STAT_L(5); for (unsigned t_ = am; t_; t_ &= t_ - 1) BVH_STAT_L(6); }
            __syncwarp();
        }
        // phase 2: one face per lane — the cheap bound first, the exactly rounded face test if it can still win
        if (gl < leaf_count) {
            BVH_STAT(2);
            BVH_STAT_L(3);
            const int s = leaf_first + gl;
            const float4 *q = reinterpret_cast<const float4 *>(tr.pre + 8 * (size_t)s);
            float l1;
            const float lbf = tri_lower_bound(q[0], q[1], p, l1);
            if (!bvh_face_skip(lbf, l1, scale, bound)) {
                BVH_STAT_L(4);
                const float *t = tr.tris + 9 * (size_t)s;
                TriHit h = tri_closest<EXACT>(p, ld3(t), ld3(t + 3), ld3(t + 6));
                sdf_best_update(best, h, tr.face_id[s]);
            }
        }
        // the gang's bound and current winner (lexicographic minimum of the lanes' bests)
        float wd = best.d2;
        int wf = best.face;
#pragma unroll
        for (int o = 1; o < GQP_GANG; o <<= 1) {
            const float od = __shfl_xor_sync(GQP_FULL, wd, o);
            const int of = __shfl_xor_sync(GQP_FULL, wf, o);
            if (od < wd || (od == wd && of < wf)) {
                wd = od;
                wf = of;
            }
        }
        bound = fminf(bound, wd);
        const bool fin = qi >= 0 && sp == 0;  // nothing pending: this search is complete
        if (__any_sync(GQP_FULL, fin)) {
            const unsigned wm = (__ballot_sync(GQP_FULL, best.face == wf) >> g0) & 0xffu;
            const int src = g0 + (wm ? __ffs((int)wm) - 1 : 0);
            SdfBest win;
            win.d2 = wd;
            win.face = wf;
            win.c = V3{__shfl_sync(GQP_FULL, best.c.x, src), __shfl_sync(GQP_FULL, best.c.y, src),
                       __shfl_sync(GQP_FULL, best.c.z, src)};
            win.normal = V3{__shfl_sync(GQP_FULL, best.normal.x, src), __shfl_sync(GQP_FULL, best.normal.y, src),
                            __shfl_sync(GQP_FULL, best.normal.z, src)};
            if (fin) {
                if (done(qi, p, tr, win)) {
                    scale = fmaxf(tr.extent, fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z))));
                    sdf_best_init(best);
                    bound = INF;
                    if (gl == 0) stk[0] = make_int2(0, 0);
                    sp = 1;
                } else {
                    qi = -1;
                }
            }
        }
    }
}
