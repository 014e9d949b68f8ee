// H4 — HandModel.cal_distance / E_pen (core/hand_model.py:875-987, core/energy.py:57-62): signed penetration depth of
// the object's surface samples into the hand's link meshes, max over links, and H5/E_joints (hand_model.py:989-1040,
// energy.py:47-54).
//
// This is the dominant stage of a grasp-step: the reference runs n_links brute-force TorchSDF launches over
// Bt*n_pts points (2500 x 3540 point-face tests per candidate for Allegro).  Two paths:
//   * culled (the engine; E_pen and its gradient only): pen_bounds -> pen_broad (a warp per candidate: row / chunk
//     bounding spheres vs link boxes) -> pen_collect (a warp per surviving (candidate, chunk): link headers, outside grids
//     and tie planes in shared memory by one TMA bulk copy; a sample is dropped for a link when it is provably outside it
//     AND not in a tie slab) -> pen_search (8-lane gangs walk the links' 8-wide trees for the queued samples, exact
//     winner; energies in 2^-40 m fixed point, gradient by float atomics) -> pen_reduce_fixed;
//   * dist-returning (HandModel.cal_distance callers): pen_mask -> hand_pen_kernel<false> (every link searched, per-sample
//     distance / argmax link / gradient stored) -> pen_reduce.
// Bound: FP32/ALU issue + L1/L2 latency.  Algorithmic HBM bytes per grasp-step: n_links*48 B of transforms + 4 B of
// energy; the surface samples (30 KB per object) and triangles stay in L2/L1.
#include "gqp_bvh.cuh"
#include "gqp_stage.cuh"
#include "graspqp_b200.h"

struct PenHdr {
    int32_t n_links, n_grids, links_off, grids_off, nodes_off, bits_off, total_bytes, planes_off;
};
struct PenLink {
    int32_t node_first, n_nodes, tri_first, n_faces;
    int32_t grid;  // index of the link's outside-grid, or -1 when the mesh is not "nice" (no culling)
    int32_t wide_first;  // first wide node of the link's 8-wide tree in GqpPenTables::d_wide
    float lo[3], hi[3];  // bounding box, slightly inflated
    int32_t plane_first, n_planes;  // tie planes of the link (gqp_bvh.cpp: gqp_mesh_tie_planes), float4 each
    int32_t anchor_first, n_anchors;  // vertices of the tie faces (float4 each, same table, after the planes)
};
// slab behind a tie plane in which the reference's fp32 first-minimum rule may return -1 for an OUTSIDE point:
// -GQP_TIE_NEG * D <= n.q - a <= GQP_TIE_POS * D, D >= distance of q to the link (fp32 spacing allows depth / distance up
// to sqrt(2^-23) = 3.45e-4, 2.43e-4 is the largest seen; the positive side covers the rounding of the sign's dot product)
#define GQP_TIE_NEG 5e-4f
#define GQP_TIE_POS 2e-6f
#define PEN_MAX_PLANES 32
#define PEN_MAX_ANCHORS 8
__device__ __forceinline__ float pen_link_reach(const PenLink &K, V3 q) {  // upper bound of |q - any point of the link|
    const V3 ctr = V3{0.5f * (K.lo[0] + K.hi[0]), 0.5f * (K.lo[1] + K.hi[1]), 0.5f * (K.lo[2] + K.hi[2])};
    const V3 h = V3{0.5f * (K.hi[0] - K.lo[0]), 0.5f * (K.hi[1] - K.lo[1]), 0.5f * (K.hi[2] - K.lo[2])};
    const V3 d = q - ctr;
    return sqrtf(dot3(d, d)) + sqrtf(dot3(h, h));
}
// true when a sphere (centre q in the link frame, radius rad; rad = 0: a point) reaches the tie slab of one of the
// link's planes: nothing in it may be culled
__device__ __forceinline__ bool pen_in_tie_slab(const PenLink &K, const float4 *planes, V3 q, float rad) {
    if (K.n_planes == 0) return false;
    float D = pen_link_reach(K, q) + rad;
    bool hit = false;
    for (int i = 0; i < K.n_planes; ++i) {
        const float4 P = planes[K.plane_first + i];
        const float hgt = P.x * q.x + P.y * q.y + P.z * q.z - P.w;
        if (hgt - rad <= GQP_TIE_POS * D && hgt + rad >= -GQP_TIE_NEG * D) hit = true;
    }
    if (!hit || rad > 0.f) return hit;
    // a point inside the coarse slab: tighten the distance bound with the nearest anchor vertex and test again
    float d2 = D * D;
    for (int i = 0; i < K.n_anchors; ++i) {
        const float4 A = planes[K.anchor_first + i];
        const float dx = q.x - A.x, dy = q.y - A.y, dz = q.z - A.z;
        d2 = fminf(d2, dx * dx + dy * dy + dz * dz);
    }
    D = sqrtf(d2) * 1.001f;
    for (int i = 0; i < K.n_planes; ++i) {
        const float4 P = planes[K.plane_first + i];
        const float hgt = P.x * q.x + P.y * q.y + P.z * q.z - P.w;
        if (hgt <= GQP_TIE_POS * D && hgt >= -GQP_TIE_NEG * D) return true;
    }
    return false;
}
struct PenGrid {
    float org[3];
    float inv_cell;
    int32_t dim[3];
    int32_t bit_first;  // first 32-bit word
};
struct GqpPenTables {
    std::vector<unsigned char> blob_host;
    unsigned char *d_blob = nullptr;
    float *d_tris = nullptr;       // (F_total, 9) leaf order, links concatenated
    int32_t *d_face_id = nullptr;  // (F_total) index inside the link's original face list
    float *d_pre = nullptr;        // (F_total, 8) pre-test records, leaf order
    GqpBvhNode *d_wide = nullptr;  // 8-wide trees of all links (gqp_bvh_widen), link l starts at wide node links[l].wide_first
    float *d_wpre = nullptr;       // patch records of the wide trees' leaf children
    float extent = 0.f;            // max |coordinate| over all link meshes
    int total_bytes = 0;
};

#ifdef GQP_EMU
#define PEN_THREADS 64
#else
#define PEN_THREADS 512
#endif
#define PEN_MIN_BLOCKS 2
#define PEN_MAX_SMEM (200 * 1024)

// true when q is provably outside the (nice) link mesh: outside its box, or in a grid cell entirely outside it
__device__ __forceinline__ bool pen_outside(const PenLink &K, const PenGrid *grids, const uint32_t *bits,
                                            const float4 *planes, V3 q) {
    if (K.grid < 0) return false;
    if (pen_in_tie_slab(K, planes, q, 0.f)) return false;
    if (q.x < K.lo[0] || q.y < K.lo[1] || q.z < K.lo[2] || q.x > K.hi[0] || q.y > K.hi[1] || q.z > K.hi[2]) return true;
    const PenGrid &G = grids[K.grid];
    float fx = (q.x - G.org[0]) * G.inv_cell, fy = (q.y - G.org[1]) * G.inv_cell, fz = (q.z - G.org[2]) * G.inv_cell;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    if (fx < 0.f || fy < 0.f || fz < 0.f || ix >= G.dim[0] || iy >= G.dim[1] || iz >= G.dim[2]) return false;
    int i = (iz * G.dim[1] + iy) * G.dim[0] + ix;
    return (bits[G.bit_first + (i >> 5)] >> (i & 31)) & 1u;
}

#define PEN_CB 8  // candidates per warp-task: amortises the sample loads and index arithmetic

// Bounding spheres of every 32-sample chunk of a point row, and of the whole row (slot n_chunks).  One block per row.
__global__ void __launch_bounds__(256)
pen_bounds_kernel(const float *__restrict__ x, int n_pts, int n_chunks, float *__restrict__ bounds) {
    __shared__ float red[8][4];
    const int row = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float *xr = x + (size_t)row * n_pts * 3;
    float *br = bounds + (size_t)row * (n_chunks + 1) * 4;
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int c = warp; c < n_chunks; c += 8) {
        const int pi = min(c * 32 + lane, n_pts - 1);
        V3 p = ld3(xr + 3 * pi);
        V3 cen = V3{warp_sum(p.x) * (1.f / 32.f), warp_sum(p.y) * (1.f / 32.f), warp_sum(p.z) * (1.f / 32.f)};
        V3 d = p - cen;
        float r = sqrtf(warp_max(dot3(d, d))) * 1.0001f + 1e-5f;
        if (lane == 0) { br[4 * c] = cen.x; br[4 * c + 1] = cen.y; br[4 * c + 2] = cen.z; br[4 * c + 3] = r; }
        if (c * 32 + lane < n_pts) { sx += p.x; sy += p.y; sz += p.z; }
    }
    sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
    if (lane == 0) { red[warp][0] = sx; red[warp][1] = sy; red[warp][2] = sz; }
    __syncthreads();
    V3 cen = V3{0, 0, 0};
    for (int w = 0; w < 8; ++w) cen = cen + V3{red[w][0], red[w][1], red[w][2]};
    cen = cen * (1.0f / n_pts);
    __syncthreads();
    float r2 = 0.f;
    for (int i = threadIdx.x; i < n_pts; i += 256) {
        V3 d = ld3(xr + 3 * i) - cen;
        r2 = fmaxf(r2, dot3(d, d));
    }
    r2 = warp_max(r2);
    if (lane == 0) red[warp][3] = r2;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = 0.f;
        for (int w = 0; w < 8; ++w) m = fmaxf(m, red[w][3]);
        br[4 * n_chunks] = cen.x; br[4 * n_chunks + 1] = cen.y; br[4 * n_chunks + 2] = cen.z;
        br[4 * n_chunks + 3] = sqrtf(m) * 1.0001f + 1e-5f;
    }
}

// Per candidate: bit l set when link l can matter for this candidate's point row, i.e. it has faces and (when culling)
// the row's bounding sphere reaches the link's box.  Exact for "nice" links (see gqp_bvh.cpp); others are always kept.
__global__ void pen_mask_kernel(const unsigned char *__restrict__ blob, const float *__restrict__ link_tf,
                                const float *__restrict__ bounds, int Bt, int batch_each, int n_chunks,
                                int per_candidate_points, int cull, uint32_t *__restrict__ masks) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(blob);
    const PenLink *links = reinterpret_cast<const PenLink *>(blob + hdr->links_off);
    const float4 *planes = reinterpret_cast<const float4 *>(blob + hdr->planes_off);
    const int L = hdr->n_links;
    const int row = per_candidate_points ? b : b / batch_each;
    const float *s = bounds + ((size_t)row * (n_chunks + 1) + n_chunks) * 4;
    const V3 cen = V3{s[0], s[1], s[2]};
    const float rad = s[3];
    uint32_t m = 0;
    for (int l = 0; l < L; ++l) {
        const PenLink &K = links[l];
        if (K.n_faces == 0) continue;  // hand_model.py:919-921: links without a mesh are skipped
        bool keep = true;
        if (cull && K.grid >= 0) {
            V3 q = t34_inv_point(t34_load(link_tf + ((size_t)b * L + l) * 12), cen);
            keep = !(q.x < K.lo[0] - rad || q.y < K.lo[1] - rad || q.z < K.lo[2] - rad || q.x > K.hi[0] + rad ||
                     q.y > K.hi[1] + rad || q.z > K.hi[2] + rad) ||
                   pen_in_tie_slab(K, planes, q, rad);
        }
        if (keep) m |= 1u << l;
    }
    masks[b] = m;
}

template <bool CULL>
__global__ void __launch_bounds__(PEN_THREADS, PEN_MIN_BLOCKS)
hand_pen_kernel(const unsigned char *__restrict__ blob, int blob_bytes, int nodes_staged, const float *__restrict__ tris,
                const int32_t *__restrict__ face_id, const float *__restrict__ pre, float extent,
                const float *__restrict__ link_tf, const float *__restrict__ x, const float *__restrict__ bounds, const uint32_t *__restrict__ masks, int n_rows, int cands_per_row,
                int n_pts, float *__restrict__ dist_out, int32_t *__restrict__ arg_link_out,
                float *__restrict__ dgrad_out, float *__restrict__ partial, float w_pen,
                float *__restrict__ grad_link_tf) {
    GQP_DYN_SMEM(unsigned char, sm);
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) mbar_init(&bar);
    __syncthreads();
    if (threadIdx.x == 0) bulk_load(sm, blob, (uint32_t)blob_bytes, &bar);
    mbar_wait(&bar, 0);
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(sm);
    const PenLink *links = reinterpret_cast<const PenLink *>(sm + hdr->links_off);
    const PenGrid *grids = reinterpret_cast<const PenGrid *>(sm + hdr->grids_off);
    // the link trees ride along in shared memory when the whole blob fits (blob_bytes == total), else they are read
    // through L1/L2 (Robotiq 2F-140: 22.7k faces, 180 KB of nodes)
    const GqpBvhNode *nodes = reinterpret_cast<const GqpBvhNode *>((nodes_staged ? sm : blob) + hdr->nodes_off);
    const uint32_t *bits = reinterpret_cast<const uint32_t *>(sm + hdr->bits_off);
    const float4 *planes = reinterpret_cast<const float4 *>(sm + hdr->planes_off);
    const int L = hdr->n_links;

    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x / 32;
    // per-warp staging of the candidate's link transforms (L x 12 floats), behind the blob
    float *sW = reinterpret_cast<float *>(sm + blob_bytes) + (size_t)(threadIdx.x >> 5) * L * 12;
    const int n_chunks = (n_pts + 31) / 32;
    const int n_cblk = (cands_per_row + PEN_CB - 1) / PEN_CB;
    const int n_tasks = n_rows * n_chunks * n_cblk;
    for (int task = blockIdx.x * warps_per_block + (threadIdx.x >> 5); task < n_tasks;
         task += gridDim.x * warps_per_block) {
        const int cb = task % n_cblk, rc = task / n_cblk;
        const int chunk = rc % n_chunks, row = rc / n_chunks;
        const int pi = chunk * 32 + lane;
        const bool valid = pi < n_pts;
        const V3 p = ld3(x + ((size_t)row * n_pts + (valid ? pi : n_pts - 1)) * 3);
        const float *sph = bounds + ((size_t)row * (n_chunks + 1) + chunk) * 4;
        const V3 cen = V3{sph[0], sph[1], sph[2]};
        const float rad = sph[3];
        const int b_end = row * cands_per_row + min(cands_per_row, (cb + 1) * PEN_CB);
        for (int b = row * cands_per_row + cb * PEN_CB; b < b_end; ++b) {
            const uint32_t mask = masks[b];
            if (mask == 0) continue;  // the whole row is outside every link of this candidate
            const float *Wb = link_tf + (size_t)b * L * 12;
            // lane l: stage link l's transform and test the chunk's sphere against the link's box
            bool keep = false;
            if ((mask >> lane) & 1u) {
                T34 W = t34_load(Wb + lane * 12);
#pragma unroll
                for (int i = 0; i < 12; ++i) sW[lane * 12 + i] = W.m[i];
                const PenLink &K = links[lane];
                keep = true;
                if (CULL && K.grid >= 0) {
                    V3 q = t34_inv_point(W, cen);
                    if ((q.x < K.lo[0] - rad || q.y < K.lo[1] - rad || q.z < K.lo[2] - rad || q.x > K.hi[0] + rad ||
                         q.y > K.hi[1] + rad || q.z > K.hi[2] + rad) &&
                        !pen_in_tie_slab(K, planes, q, rad))
                        keep = false;
                }
            }
            unsigned todo = __ballot_sync(GQP_FULL, keep);
            if (todo == 0) continue;
            __syncwarp();
            float best_d = -__int_as_float(0x7f800000);
            int best_l = -1;
            V3 best_g = V3{0, 0, 0};
            while (todo) {
                const int l = __ffs((int)todo) - 1;
                todo &= todo - 1;
                const PenLink &K = links[l];
                T34 W;
#pragma unroll
                for (int i = 0; i < 12; ++i) W.m[i] = sW[l * 12 + i];
                V3 q = t34_inv_point(W, p);
                if (CULL && pen_outside(K, grids, bits, planes, q)) continue;
                SdfBest best;
                sdf_best_init(best);
                bvh_query<false>(nodes + K.node_first, tris + (size_t)K.tri_first * 9, face_id + K.tri_first,
                                 pre + (size_t)K.tri_first * 8, extent, q, best);
                V3 dv = q - best.c;
                float sgn = dot3(dv, best.normal) >= 0.f ? 1.f : -1.f;
                float r = sqrtf(best.d2 + 1e-8f);
                float d = r * (-sgn);  // hand_model.py:975-976
                if (d > best_d) {      // :986 max over links (first maximum)
                    best_d = d;
                    best_l = l;
                    best_g = dv * (1.0f / r);
                }
            }
            if (dist_out && valid) dist_out[(size_t)b * n_pts + pi] = best_d;
            if (arg_link_out && valid) {  // saved for gqp_hand_distance_backward: argmax link and d(dist)/d(q)
                const size_t o = (size_t)b * n_pts + pi;
                const float sg = best_d > 0.f ? 1.f : -1.f;  // dist = r * (-sign): d(dist)/dq = (-sign) (q-c)/r
                arg_link_out[o] = best_l;
                st3(dgrad_out + 3 * o, best_g * sg);
            }
            float e = (valid && best_d > 0.f) ? best_d : 0.f;  // energy.py:60-61
            if (grad_link_tf && e > 0.f) {
                // d = sqrt(d2+1e-8) with sign = -1; d(d)/dq = (q-c)/r; q = R^T (p - t)
                const float *W = sW + best_l * 12;
                V3 g = best_g * w_pen;
                float *G = grad_link_tf + ((size_t)b * L + best_l) * 12;
                float pt[3] = {p.x - W[3], p.y - W[7], p.z - W[11]}, gv[3] = {g.x, g.y, g.z};
#pragma unroll
                for (int j = 0; j < 3; ++j) {
#pragma unroll
                    for (int i = 0; i < 3; ++i) atomicAdd(G + 4 * j + i, pt[j] * gv[i]);
                    atomicAdd(G + 4 * j + 3, -(W[4 * j] * gv[0] + W[4 * j + 1] * gv[1] + W[4 * j + 2] * gv[2]));
                }
            }
            if (partial && __any_sync(GQP_FULL, e > 0.f)) {  // partial was zeroed by the host wrapper
                float s = warp_sum(e);
                if (lane == 0) partial[(size_t)b * n_chunks + chunk] = s;
            }
            __syncwarp();  // sW is rewritten for the next candidate
        }
    }
}

// ---- culled path (E_pen and its gradient only): collect -> search -> reduce --------------------------------------------
// After the box / sphere / outside-grid tests only a handful of (candidate, sample) pairs per warp-task still need an
// exact search; searching them where they are found left ~2 of 32 lanes busy in the expensive part.  So the culling
// pass only EMITS the survivors — (sample id, mask of links to search) — into a global queue, and a second, dense
// kernel searches them with every lane busy.  Penetration energies are accumulated in 2^-40 m fixed point (integer
// atomics: deterministic whatever the queue order); gradients use float atomics as before.
struct PenGroup {
    uint32_t sample;  // b * n_pts + pi
    uint32_t need;    // links whose exact distance is needed
};
#define PEN_FIX_SCALE 1099511627776.0  // 2^40

// Broad phase, a warp per candidate: which links can matter for the candidate's point row at all (row sphere vs link
// box, lane = link), then, lanes = 32-sample chunks of the row, which of those links each chunk's bounding sphere
// reaches.  Chunks that reach a link become work items (candidate, chunk, link mask) for the narrow phase.  Exact for
// "nice" links (see gqp_bvh.cpp); links without an outside-grid are always kept.
struct alignas(16) PenItem {
    uint32_t b, chunk, links, row;  // row = index of the candidate's point row (saves a division per item and lane)
};
#define PEN_BROAD_WARPS 4
__global__ void __launch_bounds__(PEN_BROAD_WARPS * 32)
pen_broad_kernel(const unsigned char *__restrict__ blob, const float *__restrict__ link_tf,
                 const float *__restrict__ bounds, int Bt, int batch_each, int n_chunks, int per_candidate_points,
                 uint32_t *__restrict__ counter, PenItem *__restrict__ items) {
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * PEN_BROAD_WARPS + (threadIdx.x >> 5);
    if (b >= Bt) return;
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(blob);
    const PenLink *links = reinterpret_cast<const PenLink *>(blob + hdr->links_off);
    const float4 *planes = reinterpret_cast<const float4 *>(blob + hdr->planes_off);
    const int L = hdr->n_links;
    const int row = per_candidate_points ? b : b / batch_each;
    const float *br = bounds + (size_t)row * (n_chunks + 1) * 4;
    const float *Wb = link_tf + (size_t)b * L * 12;
    const bool tf16 = ((uintptr_t)link_tf & 15) == 0;
    bool keep = false;
    if (lane < L && links[lane].n_faces > 0) {  // hand_model.py:919-921: links without a mesh are skipped
        const PenLink &K = links[lane];
        keep = true;
        if (K.grid >= 0) {
            const float rad = br[4 * n_chunks + 3];
            const V3 q = t34_inv_point(t34_load(Wb + lane * 12), V3{br[4 * n_chunks], br[4 * n_chunks + 1], br[4 * n_chunks + 2]});
            keep = !(q.x < K.lo[0] - rad || q.y < K.lo[1] - rad || q.z < K.lo[2] - rad || q.x > K.hi[0] + rad ||
                     q.y > K.hi[1] + rad || q.z > K.hi[2] + rad) ||
                   pen_in_tie_slab(K, planes, q, rad);
        }
    }
    const uint32_t rowmask = __ballot_sync(GQP_FULL, keep);
    if (rowmask == 0) return;
    for (int c0 = 0; c0 < n_chunks; c0 += 32) {
        const int c = c0 + lane;
        uint32_t cm = 0;
        if (c < n_chunks) {
            const V3 cen = V3{br[4 * c], br[4 * c + 1], br[4 * c + 2]};  // scratch is only 4-byte aligned
            const float rad = br[4 * c + 3];
            for (uint32_t m = rowmask; m; m &= m - 1) {
                const int l = __ffs((int)m) - 1;
                const PenLink &K = links[l];
                bool hit = true;
                if (K.grid >= 0) {
                    const V3 q = t34_inv_point(tf16 ? t34_load4(Wb + l * 12) : t34_load(Wb + l * 12), cen);
                    hit = !(q.x < K.lo[0] - rad || q.y < K.lo[1] - rad || q.z < K.lo[2] - rad || q.x > K.hi[0] + rad ||
                            q.y > K.hi[1] + rad || q.z > K.hi[2] + rad) ||
                          pen_in_tie_slab(K, planes, q, rad);
                }
                if (hit) cm |= 1u << l;
            }
        }
        const unsigned em = __ballot_sync(GQP_FULL, cm != 0);
        if (em) {
            int base = 0;
            if (lane == 0) base = (int)atomicAdd(counter + 1, (unsigned)__popc(em));
            base = __shfl_sync(GQP_FULL, base, 0);
            if (cm) items[base + __popc(em & ((1u << lane) - 1u))] = PenItem{(uint32_t)b, (uint32_t)c, cm, (uint32_t)row};
        }
    }
}

// Narrow phase, a warp per work item, lane = sample: the sample is dropped for a link when it is provably outside it
// (outside the box, or in an outside cell of the link's grid); the others are queued for the exact search.
// TF16: link_tf is 16-byte aligned (128-bit loads of the transforms).
template <bool TF16>
__global__ void __launch_bounds__(PEN_THREADS)
pen_collect_kernel(const unsigned char *__restrict__ blob, int prefix_bytes, const float *__restrict__ link_tf,
                   const float *__restrict__ x, int n_pts, uint32_t *__restrict__ counter,
                   const PenItem *__restrict__ items, PenGroup *__restrict__ queue, unsigned queue_cap) {
    GQP_DYN_SMEM(unsigned char, sm);
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) mbar_init(&bar);
    __syncthreads();
    if (threadIdx.x == 0) bulk_load(sm, blob, (uint32_t)prefix_bytes, &bar);  // header, links, grids, bits: no BVH nodes
    mbar_wait(&bar, 0);
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(sm);
    const PenLink *links = reinterpret_cast<const PenLink *>(sm + hdr->links_off);
    const PenGrid *grids = reinterpret_cast<const PenGrid *>(sm + hdr->grids_off);
    const uint32_t *bits = reinterpret_cast<const uint32_t *>(sm + hdr->bits_off);
    const float4 *planes = reinterpret_cast<const float4 *>(sm + hdr->planes_off);
    const int L = hdr->n_links;
    const int lane = threadIdx.x & 31, warps_per_block = blockDim.x / 32;
    const unsigned n_items = counter[1];
    for (unsigned it = blockIdx.x * warps_per_block + (threadIdx.x >> 5); it < n_items; it += gridDim.x * warps_per_block) {
        const PenItem I = items[it];
        const int b = (int)I.b, pi = (int)I.chunk * 32 + lane;
        const size_t row = I.row;
        const bool valid = pi < n_pts;
        const V3 p = ld3(x + (row * n_pts + (valid ? pi : n_pts - 1)) * 3);
        const float *Wb = link_tf + (size_t)b * L * 12;
        uint32_t need = 0;
        for (uint32_t m = I.links; m; m &= m - 1) {
            const int l = __ffs((int)m) - 1;
            const T34 W = TF16 ? t34_load4(Wb + l * 12) : t34_load(Wb + l * 12);
            if (!pen_outside(links[l], grids, bits, planes, t34_inv_point(W, p))) need |= 1u << l;
        }
        if (!valid) need = 0;  // padding lanes of the last chunk
        const unsigned em = __ballot_sync(GQP_FULL, need != 0);
        if (em) {
            int base = 0;
            if (lane == 0) {
                base = (int)atomicAdd(counter, (unsigned)__popc(em));
                // a bounded queue that fills up sets the overflow word: the caller must treat the result as invalid
                if ((unsigned)base + (unsigned)__popc(em) > queue_cap) atomicOr(counter + 2, 1u);
            }
            base = __shfl_sync(GQP_FULL, base, 0);
            const unsigned slot = (unsigned)base + (unsigned)__popc(em & ((1u << lane) - 1u));
            if (need && slot < queue_cap) queue[slot] = PenGroup{(uint32_t)(b * n_pts + pi), need};
        }
    }
}

// Exact search of the queued samples: 8 lanes per sample over the links' 8-wide trees (gqp_bvh.cuh: bvh_query_gang),
// links in ascending order (first maximum, hand_model.py:986).  Tree nodes, patch records and faces are read through
// L1/L2 (all link meshes of a hand are ~0.3 MB), so the kernel's occupancy does not depend on the size of the tables.
#define PEN_Q_WARPS 4  // PEN_Q_QPW measured on B200 (config 2 hand_distance): 8 -> 484 us, 16 -> 518, 32 -> 598
#ifndef PEN_Q_QPW
#define PEN_Q_QPW 8
#endif  // groups per warp and round
__global__ void __launch_bounds__(PEN_Q_WARPS * 32, 8)
pen_search_kernel(const unsigned char *__restrict__ blob, const GqpBvhNode *__restrict__ wide,
                  const float *__restrict__ wpre, const float *__restrict__ tris,
                  const int32_t *__restrict__ face_id, const float *__restrict__ pre, float extent,
                  const float *__restrict__ link_tf, const float *__restrict__ x, int batch_each, int n_pts,
                  int per_candidate_points, const uint32_t *__restrict__ counter, const PenGroup *__restrict__ queue,
                  unsigned queue_cap, unsigned long long *__restrict__ acc, float w_pen,
                  float *__restrict__ grad_link_tf) {
    __shared__ int2 stk[PEN_Q_WARPS * 32 / GQP_GANG][GQP_GANG_STACK];
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(blob);
    const PenLink *links = reinterpret_cast<const PenLink *>(blob + hdr->links_off);
    const int L = hdr->n_links;
    const int n_groups = (int)min(*counter, queue_cap);
    const int gl = threadIdx.x & (GQP_GANG - 1);
    const bool tf16 = ((uintptr_t)link_tf & 15) == 0;  // 128-bit loads of the link transforms when aligned
    // per-gang state of the sample being searched (identical on the 8 lanes)
    V3 P = V3{0, 0, 0}, best_g = V3{0, 0, 0};
    uint32_t m = 0;
    int b = 0, cur_l = 0, best_l = -1;
    float best_d = 0.f;
    auto enter_link = [&](V3 &q, BvhTree &tr) {
        cur_l = __ffs((int)m) - 1;
        m &= m - 1;
        const PenLink &K = links[cur_l];
        const float *Wp = link_tf + ((size_t)b * L + cur_l) * 12;
        q = t34_inv_point(tf16 ? t34_load4(Wp) : t34_load(Wp), P);
        tr = BvhTree{wide + GQP_GANG * (size_t)K.wide_first, wpre + 8 * GQP_GANG * (size_t)K.wide_first,
                     tris + (size_t)K.tri_first * 9, face_id + K.tri_first, pre + (size_t)K.tri_first * 8, extent};
    };
    const int warp = blockIdx.x * PEN_Q_WARPS + (threadIdx.x >> 5), n_warps = gridDim.x * PEN_Q_WARPS;
    for (int base = warp * PEN_Q_QPW; base < n_groups; base += n_warps * PEN_Q_QPW) {
        bvh_query_gang<false>(
            base, min(base + PEN_Q_QPW, n_groups), stk[threadIdx.x / GQP_GANG],
            [&](int g, V3 &q, BvhTree &tr) {
                const PenGroup G = queue[g];
                b = (int)(G.sample / (unsigned)n_pts);
                const int pi = (int)(G.sample % (unsigned)n_pts);
                const size_t row = per_candidate_points ? (size_t)b : (size_t)(b / batch_each);
                P = ld3(x + (row * n_pts + pi) * 3);
                m = G.need;
                best_d = -__int_as_float(0x7f800000);
                best_l = -1;
                enter_link(q, tr);
            },
            [&](int, V3 &q, BvhTree &tr, const SdfBest &win) {
                const V3 dv = q - win.c;
                const float sgn = dot3(dv, win.normal) >= 0.f ? 1.f : -1.f;
                const float r = sqrtf(win.d2 + 1e-8f);
                const float d = r * (-sgn);  // hand_model.py:975-976
                if (d > best_d) {
                    best_d = d;
                    best_l = cur_l;
                    best_g = dv * (1.0f / r);
                }
                if (m) {
                    enter_link(q, tr);
                    return true;
                }
                if (best_d > 0.f && gl == 0) {  // energy.py:60-61
                    if (acc) atomicAdd(acc + b, (unsigned long long)__double2ll_rn((double)best_d * PEN_FIX_SCALE));
                    if (grad_link_tf) {
                        // d = sqrt(d2+1e-8) with sign = -1; d(d)/dq = (q-c)/r; q = R^T (p - t)
                        const size_t wl = ((size_t)b * L + best_l) * 12;
                        const float *W = link_tf + wl;
                        float *Gm = grad_link_tf + wl;
                        const V3 gg = best_g * w_pen;
                        const float pt[3] = {P.x - W[3], P.y - W[7], P.z - W[11]}, gv[3] = {gg.x, gg.y, gg.z};
#pragma unroll
                        for (int r2 = 0; r2 < 3; ++r2) {
#pragma unroll
                            for (int i = 0; i < 3; ++i) atomicAdd(Gm + 4 * r2 + i, pt[r2] * gv[i]);
                            atomicAdd(Gm + 4 * r2 + 3,
                                      -(W[4 * r2] * gv[0] + W[4 * r2 + 1] * gv[1] + W[4 * r2 + 2] * gv[2]));
                        }
                    }
                }
                return false;
            });
    }
}

__global__ void pen_reduce_fixed_kernel(const unsigned long long *__restrict__ acc, int Bt, float *__restrict__ e_pen) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    e_pen[b] = (float)((double)acc[b] * (1.0 / PEN_FIX_SCALE));
}

__global__ void pen_reduce_kernel(const float *__restrict__ partial, int Bt, int n_chunks, float *__restrict__ e_pen) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    float s = 0.f;
    for (int c = 0; c < n_chunks; ++c) s += partial[(size_t)b * n_chunks + c];
    e_pen[b] = s;
}

// ---- H5 + E_joints: a thread per candidate --------------------------------------------------------------------------
#define GQP_MAX_SPHERES 64
__global__ void __launch_bounds__(128)
self_pen_joints_kernel(const GqpHandDev *__restrict__ hand, const float *__restrict__ spheres,
                       const int32_t *__restrict__ sphere_link, const float *__restrict__ link_tf,
                       const float *__restrict__ pose, int Bt, float *__restrict__ e_spen, float *__restrict__ e_joints,
                       float w_spen, float w_joints, float *__restrict__ grad_link_tf,
                       float *__restrict__ grad_pose_joints) {
    __shared__ float s_sph[GQP_MAX_SPHERES * 4];
    __shared__ int32_t s_link[GQP_MAX_SPHERES];
    __shared__ float s_lo[GQP_MAX_DOFS], s_hi[GQP_MAX_DOFS];
    const int S = hand->n_spheres, L = hand->n_links, dof = hand->n_dofs;
    for (int i = threadIdx.x; i < S * 4; i += blockDim.x) s_sph[i] = spheres[i];
    for (int i = threadIdx.x; i < S; i += blockDim.x) s_link[i] = sphere_link[i];
    for (int i = threadIdx.x; i < dof; i += blockDim.x) {
        s_lo[i] = hand->joints_lower[i];
        s_hi[i] = hand->joints_upper[i];
    }
    __syncthreads();
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bt) return;
    const int D = 9 + dof;
    // E_joints (energy.py:47-54)
    float ej = 0.f;
    for (int j = 0; j < dof; ++j) {
        float th = pose[(size_t)b * D + 9 + j];
        float g = 0.f;
        if (th > s_hi[j]) { ej += th - s_hi[j]; g = 1.f; }
        if (th < s_lo[j]) { ej += s_lo[j] - th; g = -1.f; }
        if (grad_pose_joints) grad_pose_joints[(size_t)b * D + 9 + j] = g * w_joints;
    }
    if (grad_pose_joints)
        for (int j = 0; j < 9; ++j) grad_pose_joints[(size_t)b * D + j] = 0.f;
    if (e_joints) e_joints[b] = ej;
    // self penetration (hand_model.py:989-1040)
    float wp[GQP_MAX_SPHERES][3];
    for (int i = 0; i < S; ++i) {
        T34 W = t34_load(link_tf + ((size_t)b * L + s_link[i]) * 12);
        V3 c = t34_point(W, V3{s_sph[4 * i], s_sph[4 * i + 1], s_sph[4 * i + 2]});
        wp[i][0] = c.x; wp[i][1] = c.y; wp[i][2] = c.z;
    }
    float es = 0.f;
    int i0 = 0;
    while (i0 < S) {
        int i1 = i0;
        while (i1 < S && s_link[i1] == s_link[i0]) ++i1;  // spheres of one link: [i0, i1)
        if (i1 >= S) break;                               // the last link has nothing after it (:1013)
        float mn = __int_as_float(0x7f800000);
        int ma = -1, mb = -1;
        for (int a = i0; a < i1; ++a)
            for (int c = i1; c < S; ++c) {
                float dx = wp[a][0] - wp[c][0] + 1e-13f, dy = wp[a][1] - wp[c][1] + 1e-13f,
                      dz = wp[a][2] - wp[c][2] + 1e-13f;
                float pen = sqrtf(dx * dx + dy * dy + dz * dz) - (s_sph[4 * a + 3] + s_sph[4 * c + 3]);
                if (pen < mn) { mn = pen; ma = a; mb = c; }
            }
        if (mn < 0.f) {
            es += -mn;
            if (grad_link_tf) {
                float dx = wp[ma][0] - wp[mb][0] + 1e-13f, dy = wp[ma][1] - wp[mb][1] + 1e-13f,
                      dz = wp[ma][2] - wp[mb][2] + 1e-13f;
                float inv = -w_spen / sqrtf(dx * dx + dy * dy + dz * dz);  // d(-dis)/da = -(a-b)/dis
                float g[3] = {dx * inv, dy * inv, dz * inv};
                for (int side = 0; side < 2; ++side) {
                    int k = side == 0 ? ma : mb;
                    float sg = side == 0 ? 1.f : -1.f;
                    float *G = grad_link_tf + ((size_t)b * L + s_link[k]) * 12;
                    for (int r = 0; r < 3; ++r) {
                        for (int q = 0; q < 3; ++q) atomicAdd(G + 4 * r + q, sg * g[r] * s_sph[4 * k + q]);
                        atomicAdd(G + 4 * r + 3, sg * g[r]);
                    }
                }
            }
        }
        i0 = i1;
    }
    if (e_spen) e_spen[b] = es;
}

// ---- host side ------------------------------------------------------------------------------------------------------
static void pad16(std::vector<unsigned char> &v) { v.resize((v.size() + 15) / 16 * 16, 0); }

int gqp_pen_tables_create(GqpHand *h, const float *face_verts_host) {
    const int L = h->host.n_links;
    GqpPenTables *T = new GqpPenTables();
    h->pen = T;
    std::vector<PenLink> links(L);
    std::vector<PenGrid> grids;
    std::vector<GqpBvhNode> nodes;
    std::vector<uint32_t> bits;
    std::vector<float> tris;
    std::vector<int32_t> fid;
    std::vector<float> pre;
    std::vector<GqpBvhNode> wide;
    std::vector<float> wpre;
    std::vector<float> planes;
    for (int l = 0; l < L; ++l) {
        int f0 = h->face_offset[l], F = h->face_offset[l + 1] - f0;
        PenLink K{};
        K.node_first = (int)nodes.size();
        K.wide_first = (int)(wide.size() / 8);
        K.tri_first = (int)(tris.size() / 9);
        K.n_faces = F;
        K.grid = -1;
        if (F > 0) {
            GqpBvhHost bvh;
            gqp_bvh_build(face_verts_host + 9 * (size_t)f0, F, 8, bvh);
            if (bvh.max_depth > GQP_BVH_STACK - 2)
                return gqp_set_error(GQP_ERR_LIMIT, "hand link %d: BVH too deep (%d)", l, bvh.max_depth);
            gqp_bvh_widen(bvh);
            if (7 * bvh.wide_depth + 1 > GQP_GANG_STACK)
                return gqp_set_error(GQP_ERR_LIMIT, "hand link %d: wide BVH too deep (%d)", l, bvh.wide_depth);
            wide.insert(wide.end(), bvh.wide.begin(), bvh.wide.end());
            wpre.insert(wpre.end(), bvh.wide_pre.begin(), bvh.wide_pre.end());
            K.n_nodes = (int)bvh.nodes.size();
            nodes.insert(nodes.end(), bvh.nodes.begin(), bvh.nodes.end());
            tris.insert(tris.end(), bvh.tris.begin(), bvh.tris.end());
            fid.insert(fid.end(), bvh.face_id.begin(), bvh.face_id.end());
            pre.insert(pre.end(), bvh.pre.begin(), bvh.pre.end());
            T->extent = std::max(T->extent, bvh.extent);
            const GqpBvhNode &root = bvh.nodes[0];
            for (int k = 0; k < 3; ++k) {
                K.lo[k] = root.lo[k] - 2e-5f;
                K.hi[k] = root.hi[k] + 2e-5f;
            }
            std::vector<float> tie, anchors;
            if (gqp_mesh_is_nice(face_verts_host + 9 * (size_t)f0, F) &&
                gqp_mesh_tie_planes(face_verts_host + 9 * (size_t)f0, F, PEN_MAX_PLANES, tie, anchors, PEN_MAX_ANCHORS)) {
                K.plane_first = (int)(planes.size() / 4);
                K.n_planes = (int)(tie.size() / 4);
                planes.insert(planes.end(), tie.begin(), tie.end());
                K.anchor_first = (int)(planes.size() / 4);
                K.n_anchors = (int)(anchors.size() / 4);
                planes.insert(planes.end(), anchors.begin(), anchors.end());
                PenGrid G{};
                float cell = 0.002f;
                int dim[3];
                for (;;) {
                    bool ok = true;
                    for (int k = 0; k < 3; ++k) {
                        dim[k] = (int)std::ceil((K.hi[k] - K.lo[k]) / cell) + 1;
                        if (dim[k] > 48) ok = false;
                    }
                    if (ok) break;
                    cell *= 1.1f;
                }
                G.inv_cell = 1.0f / cell;
                for (int k = 0; k < 3; ++k) {
                    G.dim[k] = dim[k];
                    G.org[k] = 0.5f * (K.lo[k] + K.hi[k]) - 0.5f * dim[k] * (1.0f / G.inv_cell);
                }
                std::vector<uint32_t> gb;
                gqp_outside_grid_build(bvh, G.org, 1.0f / G.inv_cell, G.dim, gb);
                G.bit_first = (int)bits.size();
                bits.insert(bits.end(), gb.begin(), gb.end());
                K.grid = (int)grids.size();
                grids.push_back(G);
            }
        }
        links[l] = K;
    }
    PenHdr hdr{};
    hdr.n_links = L;
    hdr.n_grids = (int)grids.size();
    std::vector<unsigned char> &B = T->blob_host;
    B.resize(sizeof(PenHdr));
    auto append = [&](const void *p, size_t n) {
        pad16(B);
        int off = (int)B.size();
        B.insert(B.end(), (const unsigned char *)p, (const unsigned char *)p + n);
        return off;
    };
    hdr.links_off = append(links.data(), links.size() * sizeof(PenLink));
    hdr.grids_off = append(grids.data(), grids.size() * sizeof(PenGrid));
    hdr.bits_off = append(bits.data(), bits.size() * sizeof(uint32_t));
    planes.resize(planes.size() + 4, 0.f);  // never empty: the kernels form the pointer unconditionally
    hdr.planes_off = append(planes.data(), planes.size() * sizeof(float));
    hdr.nodes_off = append(nodes.data(), nodes.size() * sizeof(GqpBvhNode));  // last: the collect kernel stops before it
    pad16(B);
    hdr.total_bytes = (int)B.size();
    std::memcpy(B.data(), &hdr, sizeof(hdr));
    if (hdr.nodes_off > PEN_MAX_SMEM)  // links, grids and outside-grid bits must fit; the trees may stay in global memory
        return gqp_set_error(GQP_ERR_LIMIT, "hand penetration tables need %d B of shared memory (limit %d)",
                             hdr.nodes_off, PEN_MAX_SMEM);
    T->total_bytes = hdr.total_bytes;
    T->d_blob = (unsigned char *)gqp_dev_alloc_upload(B.data(), B.size());
    T->d_tris = (float *)gqp_dev_alloc_upload(tris.data(), tris.size() * sizeof(float));
    T->d_face_id = (int32_t *)gqp_dev_alloc_upload(fid.data(), fid.size() * sizeof(int32_t));
    T->d_pre = (float *)gqp_dev_alloc_upload(pre.data(), pre.size() * sizeof(float));
    T->d_wide = (GqpBvhNode *)gqp_dev_alloc_upload(wide.data(), wide.size() * sizeof(GqpBvhNode));
    T->d_wpre = (float *)gqp_dev_alloc_upload(wpre.data(), wpre.size() * sizeof(float));
    if (!T->d_blob || !T->d_tris || !T->d_face_id || !T->d_pre || !T->d_wide || !T->d_wpre)
        return gqp_set_error(GQP_ERR_CUDA, "pen tables: allocation failed");
#ifndef GQP_EMU
    cudaFuncSetAttribute(pen_collect_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PEN_MAX_SMEM + 24 * 1024);
    cudaFuncSetAttribute(pen_collect_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PEN_MAX_SMEM + 24 * 1024);
    cudaFuncSetAttribute(hand_pen_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PEN_MAX_SMEM + 24 * 1024);
#endif
    return GQP_OK;
}

void gqp_pen_tables_destroy(GqpHand *h) {
    if (!h || !h->pen) return;
    gqp_dev_free(h->pen->d_blob); gqp_dev_free(h->pen->d_tris); gqp_dev_free(h->pen->d_face_id);
    gqp_dev_free(h->pen->d_pre); gqp_dev_free(h->pen->d_wide); gqp_dev_free(h->pen->d_wpre);
    delete h->pen;
    h->pen = nullptr;
}

static int pen_grid_size(int blob_bytes, long long n_tasks, int max_per_sm) {
#ifndef GQP_EMU
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int per_sm = (int)std::max(1, std::min(max_per_sm, (int)((226 * 1024) / (blob_bytes + 1024))));
    long long want = (n_tasks + PEN_THREADS / 32 - 1) / (PEN_THREADS / 32);
    return (int)std::max(1LL, std::min<long long>(want, (long long)sms * per_sm));
#else
    (void)blob_bytes;
    (void)max_per_sm;
    return (int)std::max(1LL, std::min<long long>(2, n_tasks));
#endif
}

// the search kernel is persistent over a queue whose length only the device knows: fill the GPU
static int pen_search_grid() {
#ifndef GQP_EMU
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    static int per_sm = 0;  // tuning knob (profiles/r2_notes.md): resident CTAs per SM the search may hold
    if (per_sm == 0) {
        const char *e = getenv("GQP_PEN_SEARCH_CTAS");
        per_sm = e ? std::max(1, std::min(8, atoi(e))) : 8;
    }
    return sms * per_sm;
#else
    return 2;
#endif
}

// queue capacity of the culled path: samples per candidate on average (0 = n_pts: can never fill up)
static unsigned long long pen_queue_cap(unsigned long long Bt, int n_pts, int queue_per_candidate) {
    const int q = queue_per_candidate > 0 ? std::min(queue_per_candidate, n_pts) : n_pts;
    return Bt * (unsigned long long)q;
}
static int hand_distance_impl(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each,
                              int n_pts, int per_candidate_points, float *dist, int32_t *arg_link, float *dgrad,
                              float *e_pen, float *scratch, float w_pen, float *grad_link_tf, void *stream,
                              int queue_per_candidate = 0) {
    if (!h || n_obj < 0 || batch_each < 0 || n_pts <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_distance: bad argument");
    const int Bt = n_obj * batch_each;
    if (Bt == 0) return GQP_OK;
    if (!link_tf || !x || !scratch) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_distance: null pointer (scratch is required)");
    cudaStream_t st = (cudaStream_t)stream;
    const int n_chunks = (n_pts + 31) / 32;
    const int n_rows = per_candidate_points ? Bt : n_obj, cands_per_row = per_candidate_points ? 1 : batch_each;
    const long long n_tasks = (long long)n_rows * n_chunks * ((cands_per_row + PEN_CB - 1) / PEN_CB);
    if (n_tasks > 0x7fffffffLL) return gqp_set_error(GQP_ERR_LIMIT, "gqp_hand_distance: too many point chunks");
    if ((long long)Bt * n_pts > 0x7fffffffLL)  // queued sample ids are 32-bit (b * n_pts + sample)
        return gqp_set_error(GQP_ERR_LIMIT, "gqp_hand_distance: candidates x samples exceeds 2^31 (%d x %d)", Bt, n_pts);
    const GqpPenTables *T = h->pen;
    // scratch: [Bt*n_chunks partial sums][n_rows*(n_chunks+1)*4 bounding spheres][Bt link masks]
    float *partial = scratch;
    float *bounds = scratch + (size_t)Bt * n_chunks;
    uint32_t *masks = reinterpret_cast<uint32_t *>(bounds + (size_t)n_rows * (n_chunks + 1) * 4);
    const int cull = dist ? 0 : 1;  // with dist wanted, every link's true distance is needed
    int rc = GQP_OK;
    GQP_LAUNCH(pen_bounds_kernel, n_rows, 256, 0, st, x, n_pts, n_chunks, bounds);
    GQP_CHECK_LAUNCH("gqp_hand_distance/bounds");
    if (cull) {
        // [counters 4 words][fixed-point energy accumulators Bt x u64][queue Bt*n_pts groups][items Bt*n_chunks]
        uint32_t *tail = masks + Bt;
        tail += ((uintptr_t)tail & 7) ? 1 : 0;
        uint32_t *counter = tail;  // [0] groups queued for the search, [1] work items of the narrow phase
        unsigned long long *acc = reinterpret_cast<unsigned long long *>(tail + 4);
        PenGroup *queue = reinterpret_cast<PenGroup *>(acc + Bt);
        const unsigned long long qcap = pen_queue_cap((unsigned long long)Bt, n_pts, queue_per_candidate);
        PenItem *items = reinterpret_cast<PenItem *>(queue + qcap);
        items = reinterpret_cast<PenItem *>(((uintptr_t)items + 15) & ~(uintptr_t)15);
        rc = gqp_memset_zero(tail, 16 + sizeof(unsigned long long) * (size_t)Bt, stream);
        if (rc != GQP_OK) return rc;
        GQP_LAUNCH(pen_broad_kernel, (Bt + PEN_BROAD_WARPS - 1) / PEN_BROAD_WARPS, PEN_BROAD_WARPS * 32, 0, st, T->d_blob,
                   link_tf, bounds, Bt, batch_each, n_chunks, per_candidate_points, counter, items);
        GQP_CHECK_LAUNCH("gqp_hand_distance/broad");
        const PenHdr *hdr = reinterpret_cast<const PenHdr *>(T->blob_host.data());
        const int prefix = hdr->nodes_off;  // header, links, grids, outside-grid bits
        int grid1 = pen_grid_size(prefix, (long long)Bt * n_chunks, 3);
        if (((uintptr_t)link_tf & 15) == 0) {
            GQP_LAUNCH(pen_collect_kernel<true>, grid1, PEN_THREADS, (size_t)prefix, st, T->d_blob, prefix, link_tf, x,
                       n_pts, counter, items, queue, (unsigned)qcap);
        } else {
            GQP_LAUNCH(pen_collect_kernel<false>, grid1, PEN_THREADS, (size_t)prefix, st, T->d_blob, prefix, link_tf, x,
                       n_pts, counter, items, queue, (unsigned)qcap);
        }
        GQP_CHECK_LAUNCH("gqp_hand_distance/collect");
        GQP_LAUNCH(pen_search_kernel, pen_search_grid(), PEN_Q_WARPS * 32, 0, st, T->d_blob, T->d_wide, T->d_wpre, T->d_tris,
                   T->d_face_id, T->d_pre, T->extent, link_tf, x, batch_each, n_pts, per_candidate_points, counter, queue,
                   (unsigned)qcap, e_pen ? acc : nullptr, w_pen, grad_link_tf);
        GQP_CHECK_LAUNCH("gqp_hand_distance/search");
        if (e_pen) {
            GQP_LAUNCH(pen_reduce_fixed_kernel, (Bt + 127) / 128, 128, 0, st, acc, Bt, e_pen);
            GQP_CHECK_LAUNCH("gqp_hand_distance/reduce");
        }
        return GQP_OK;
    }
    GQP_LAUNCH(pen_mask_kernel, (Bt + 127) / 128, 128, 0, st, T->d_blob, link_tf, bounds, Bt, batch_each, n_chunks,
               per_candidate_points, cull, masks);
    GQP_CHECK_LAUNCH("gqp_hand_distance/mask");
    rc = gqp_memset_zero(partial, sizeof(float) * (size_t)Bt * n_chunks, stream);
    if (rc != GQP_OK) return rc;
    const int nodes_staged = T->total_bytes <= PEN_MAX_SMEM;
    const int staged = nodes_staged ? T->total_bytes : reinterpret_cast<const PenHdr *>(T->blob_host.data())->nodes_off;
    const size_t smem = (size_t)staged + (size_t)(PEN_THREADS / 32) * (h->host.n_links * 48);
    int grid = pen_grid_size((int)smem, n_tasks, PEN_MIN_BLOCKS);
    GQP_LAUNCH(hand_pen_kernel<false>, grid, PEN_THREADS, smem, st, T->d_blob, staged, nodes_staged, T->d_tris, T->d_face_id,
               T->d_pre, T->extent, link_tf, x, bounds, masks, n_rows, cands_per_row, n_pts, dist, arg_link, dgrad,
               partial, w_pen, grad_link_tf);
    GQP_CHECK_LAUNCH("gqp_hand_distance");
    if (e_pen) {
        GQP_LAUNCH(pen_reduce_kernel, (Bt + 127) / 128, 128, 0, st, partial, Bt, n_chunks, e_pen);
        GQP_CHECK_LAUNCH("gqp_hand_distance/reduce");
    }
    return GQP_OK;
}

extern "C" int gqp_hand_distance(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each,
                                 int n_pts, int per_candidate_points, float *dist, float *e_pen, float *scratch,
                                 float w_pen, float *grad_link_tf, void *stream) {
    return hand_distance_impl(h, link_tf, x, n_obj, batch_each, n_pts, per_candidate_points, dist, nullptr, nullptr,
                              e_pen, scratch, w_pen, grad_link_tf, stream);
}

// cal_distance with what its backward needs: argmax link and d(dist)/d(link-frame point) per sample
extern "C" int gqp_hand_distance_full(const GqpHand *h, const float *link_tf, const float *x, int n_obj,
                                      int batch_each, int n_pts, int per_candidate_points, float *dist,
                                      int32_t *arg_link, float *dgrad, float *scratch, void *stream) {
    if (!dist || !arg_link || !dgrad) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_distance_full: null output");
    return hand_distance_impl(h, link_tf, x, n_obj, batch_each, n_pts, per_candidate_points, dist, arg_link, dgrad,
                              nullptr, scratch, 0.f, nullptr, stream);
}

// d(loss)/d(link_tf) += sum_p grad_dist[b,p] * d(dist[b,p])/d(link_tf[b, arg_link])   (q = R^T (x - t))
__global__ void hand_distance_backward_kernel(const float *__restrict__ link_tf, const float *__restrict__ x,
                                              const float *__restrict__ grad_dist, const int32_t *__restrict__ arg_link,
                                              const float *__restrict__ dgrad, int Bt, int batch_each, int n_pts,
                                              int per_candidate_points, int L, float *__restrict__ grad_link_tf) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)Bt * n_pts) return;
    const float g0 = grad_dist[i];
    const int l = arg_link[i];
    if (g0 == 0.f || l < 0) return;
    const int b = (int)(i / n_pts), pi = (int)(i % n_pts);
    const size_t row = per_candidate_points ? (size_t)b : (size_t)(b / batch_each);
    const V3 p = ld3(x + (row * n_pts + pi) * 3);
    const float *W = link_tf + ((size_t)b * L + l) * 12;
    float *G = grad_link_tf + ((size_t)b * L + l) * 12;
    const float gv[3] = {dgrad[3 * i] * g0, dgrad[3 * i + 1] * g0, dgrad[3 * i + 2] * g0};
    const float pt[3] = {p.x - W[3], p.y - W[7], p.z - W[11]};
    for (int j = 0; j < 3; ++j) {
        for (int k = 0; k < 3; ++k) atomicAdd(G + 4 * j + k, pt[j] * gv[k]);
        atomicAdd(G + 4 * j + 3, -(W[4 * j] * gv[0] + W[4 * j + 1] * gv[1] + W[4 * j + 2] * gv[2]));
    }
}
extern "C" int gqp_hand_distance_backward(const GqpHand *h, const float *link_tf, const float *x, int n_obj,
                                          int batch_each, int n_pts, int per_candidate_points, const float *grad_dist,
                                          const int32_t *arg_link, const float *dgrad, float *grad_link_tf,
                                          void *stream) {
    if (!h || n_obj < 0 || batch_each < 0 || n_pts <= 0) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_distance_backward: bad argument");
    const long long n = (long long)n_obj * batch_each * n_pts;
    if (n == 0) return GQP_OK;
    if (!link_tf || !x || !grad_dist || !arg_link || !dgrad || !grad_link_tf)
        return gqp_set_error(GQP_ERR_ARG, "gqp_hand_distance_backward: null pointer");
    GQP_LAUNCH(hand_distance_backward_kernel, (unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream, link_tf, x,
               grad_dist, arg_link, dgrad, n_obj * batch_each, batch_each, n_pts, per_candidate_points,
               h->host.n_links, grad_link_tf);
    GQP_CHECK_LAUNCH("gqp_hand_distance_backward");
    return GQP_OK;
}

extern "C" int gqp_self_pen_joints(const GqpHand *h, const float *link_tf, const float *pose, int Bt, float *e_spen,
                                   float *e_joints, float w_spen, float w_joints, float *grad_link_tf,
                                   float *grad_pose_joints, void *stream) {
    if (!h || Bt < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_self_pen_joints: bad argument");
    if (Bt == 0) return GQP_OK;
    if (!link_tf || !pose) return gqp_set_error(GQP_ERR_ARG, "gqp_self_pen_joints: null pointer");
    if (h->host.n_spheres > GQP_MAX_SPHERES) return gqp_set_error(GQP_ERR_LIMIT, "too many penetration spheres");
    GQP_LAUNCH(self_pen_joints_kernel, (Bt + 127) / 128, 128, 0, (cudaStream_t)stream, h->d_dev, h->d_spheres,
               h->d_sphere_link, link_tf, pose, Bt, e_spen, e_joints, w_spen, w_joints, grad_link_tf,
               grad_pose_joints);
    GQP_CHECK_LAUNCH("gqp_self_pen_joints");
    return GQP_OK;
}

// diagnostics for DESIGN.md / tests: {blob bytes, BVH nodes, grid cells, cells proven positive, faces, grids}
extern "C" int gqp_hand_pen_info(const GqpHand *h, int32_t *info_host) {
    if (!h || !h->pen || !info_host) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_pen_info: bad argument");
    const std::vector<unsigned char> &B = h->pen->blob_host;
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(B.data());
    const PenGrid *grids = reinterpret_cast<const PenGrid *>(B.data() + hdr->grids_off);
    const uint32_t *bits = reinterpret_cast<const uint32_t *>(B.data() + hdr->bits_off);
    long long cells = 0, set = 0;
    for (int g = 0; g < hdr->n_grids; ++g) {
        long long n = (long long)grids[g].dim[0] * grids[g].dim[1] * grids[g].dim[2];
        cells += n;
        for (long long i = 0; i < n; ++i) set += (bits[grids[g].bit_first + (i >> 5)] >> (i & 31)) & 1u;
    }
    info_host[0] = hdr->total_bytes;
    info_host[1] = (hdr->total_bytes - hdr->nodes_off) / (int)sizeof(GqpBvhNode);
    info_host[2] = (int32_t)cells;
    info_host[3] = (int32_t)set;
    info_host[4] = h->face_offset.back();
    info_host[5] = hdr->n_grids;
    return GQP_OK;
}

// diagnostics for the culling-proof tests: the outside grid of one link, as the kernels use it
extern "C" int gqp_hand_link_grid(const GqpHand *h, int link, float *geom_host, int32_t *dim_host, uint32_t *bits_host,
                                  int max_words) {
    if (!h || !h->pen || !geom_host || !dim_host) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_grid: bad argument");
    const std::vector<unsigned char> &B = h->pen->blob_host;
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(B.data());
    if (link < 0 || link >= hdr->n_links) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_grid: no such link");
    const PenLink &K = reinterpret_cast<const PenLink *>(B.data() + hdr->links_off)[link];
    for (int k = 0; k < 3; ++k) {
        geom_host[4 + k] = K.lo[k];
        geom_host[7 + k] = K.hi[k];
    }
    dim_host[0] = dim_host[1] = dim_host[2] = 0;
    geom_host[0] = geom_host[1] = geom_host[2] = geom_host[3] = 0.f;
    if (K.grid < 0) return GQP_OK;
    const PenGrid &G = reinterpret_cast<const PenGrid *>(B.data() + hdr->grids_off)[K.grid];
    const uint32_t *bits = reinterpret_cast<const uint32_t *>(B.data() + hdr->bits_off) + G.bit_first;
    for (int k = 0; k < 3; ++k) {
        geom_host[k] = G.org[k];
        dim_host[k] = G.dim[k];
    }
    geom_host[3] = G.inv_cell;
    const long long words = ((long long)G.dim[0] * G.dim[1] * G.dim[2] + 31) / 32;
    if (bits_host) {
        if (max_words < words) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_grid: bits_host too small");
        std::memcpy(bits_host, bits, (size_t)words * 4);
    }
    return GQP_OK;
}

// diagnostics: word offset, inside the scratch buffer of gqp_hand_distance, of the culled path's two counters
// {samples queued for the exact search, (candidate, chunk) work items of the narrow phase} after a call
extern "C" uint64_t gqp_hand_distance_counter_offset(int n_obj, int batch_each, int n_pts, int per_candidate_points,
                                                     const void *scratch) {
    const uint64_t Bt = (uint64_t)n_obj * batch_each, n_chunks = (n_pts + 31) / 32;
    const uint64_t n_rows = per_candidate_points ? Bt : (uint64_t)n_obj;
    uint64_t off = Bt * n_chunks + n_rows * (n_chunks + 1) * 4 + Bt;  // partial sums, bounds, masks (hand_distance_impl)
    if (((uintptr_t)scratch + 4 * off) & 7) off += 1;
    return off;
}

// diagnostics: the link's tie planes (4 floats each) and the slab constants {GQP_TIE_NEG, GQP_TIE_POS}; returns the count
extern "C" int gqp_hand_link_planes(const GqpHand *h, int link, float *planes_host, int max_planes, float *slab_host) {
    if (!h || !h->pen) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_planes: bad argument");
    const std::vector<unsigned char> &B = h->pen->blob_host;
    const PenHdr *hdr = reinterpret_cast<const PenHdr *>(B.data());
    if (link < 0 || link >= hdr->n_links) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_planes: no such link");
    const PenLink &K = reinterpret_cast<const PenLink *>(B.data() + hdr->links_off)[link];
    if (slab_host) {
        slab_host[0] = GQP_TIE_NEG;
        slab_host[1] = GQP_TIE_POS;
    }
    if (planes_host) {
        if (max_planes < K.n_planes + K.n_anchors) return gqp_set_error(GQP_ERR_ARG, "gqp_hand_link_planes: planes_host too small");
        std::memcpy(planes_host, B.data() + hdr->planes_off + 16 * (size_t)K.plane_first,
                    16 * (size_t)(K.n_planes + K.n_anchors));  // the anchors follow the planes in the table
    }
    if (slab_host) slab_host[2] = (float)K.n_anchors;
    return K.n_planes;
}

extern "C" uint64_t gqp_hand_distance_scratch_floats_bounded(int n_obj, int batch_each, int n_pts,
                                                             int per_candidate_points, int queue_per_candidate) {
    const uint64_t Bt = (uint64_t)n_obj * batch_each, n_chunks = (n_pts + 31) / 32;
    const uint64_t n_rows = per_candidate_points ? Bt : (uint64_t)n_obj;
    // + culled path: counters (4 words + alignment), fixed-point accumulators (2 words each), queue (2 words per
    // queued sample), work items of the narrow phase (4 words per candidate and chunk, 16-byte aligned)
    return Bt * n_chunks + n_rows * (n_chunks + 1) * 4 + Bt + 16 + 8 + 2 * Bt +
           2 * pen_queue_cap(Bt, n_pts, queue_per_candidate) + 4 * Bt * n_chunks + 4;
}
extern "C" uint64_t gqp_hand_distance_scratch_floats(int n_obj, int batch_each, int n_pts, int per_candidate_points) {
    return gqp_hand_distance_scratch_floats_bounded(n_obj, batch_each, n_pts, per_candidate_points, 0);
}
extern "C" int gqp_hand_distance_bounded(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each,
                                         int n_pts, int per_candidate_points, float *e_pen, float *scratch, float w_pen,
                                         float *grad_link_tf, int queue_per_candidate, void *stream) {
    return hand_distance_impl(h, link_tf, x, n_obj, batch_each, n_pts, per_candidate_points, nullptr, nullptr, nullptr,
                              e_pen, scratch, w_pen, grad_link_tf, stream, queue_per_candidate);
}
