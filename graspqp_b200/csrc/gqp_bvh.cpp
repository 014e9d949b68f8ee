// Host-side builders: BVH over a triangle mesh and the far-field sign grids (see gqp_hand.h).  Runs once per mesh at
// *_create time; nothing here is on the per-iteration path.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

#include "gqp_hand.h"

namespace {
struct D3 {
    double x, y, z;
};
inline D3 operator-(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline D3 operator+(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline D3 operator*(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 ldv(const float *p) { return {p[0], p[1], p[2]}; }

// exact (double) distance point -> triangle, Ericson's region walk
double tri_dist(D3 p, D3 a, D3 b, D3 c) {
    D3 ab = b - a, ac = c - a, ap = p - a;
    double d1 = dot(ab, ap), d2 = dot(ac, ap);
    D3 q;
    if (d1 <= 0 && d2 <= 0) q = a;
    else {
        D3 bp = p - b;
        double d3 = dot(ab, bp), d4 = dot(ac, bp);
        if (d3 >= 0 && d4 <= d3) q = b;
        else {
            double vc = d1 * d4 - d3 * d2;
            if (vc <= 0 && d1 >= 0 && d3 <= 0) q = a + ab * (d1 / (d1 - d3));
            else {
                D3 cp = p - c;
                double d5 = dot(ab, cp), d6 = dot(ac, cp);
                if (d6 >= 0 && d5 <= d6) q = c;
                else {
                    double vb = d5 * d2 - d1 * d6;
                    if (vb <= 0 && d2 >= 0 && d6 <= 0) q = a + ac * (d2 / (d2 - d6));
                    else {
                        double va = d3 * d6 - d5 * d4;
                        if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0)
                            q = b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
                        else {
                            double den = 1.0 / (va + vb + vc);
                            q = a + ab * (vb * den) + ac * (vc * den);
                        }
                    }
                }
            }
        }
    }
    D3 d = p - q;
    return std::sqrt(dot(d, d));
}

double box_dist(const GqpBvhNode &n, D3 p) {
    double dx = std::max({0.0, (double)n.lo[0] - p.x, p.x - (double)n.hi[0]});
    double dy = std::max({0.0, (double)n.lo[1] - p.y, p.y - (double)n.hi[1]});
    double dz = std::max({0.0, (double)n.lo[2] - p.z, p.z - (double)n.hi[2]});
    return std::sqrt(dx * dx + dy * dy + dz * dz);
}

struct Builder {
    const float *fv;
    int leaf_size;
    std::vector<int> order;
    std::vector<float> cen;  // (F,3)
    GqpBvhHost *out;
    float pad;

    int build(int lo, int hi, int depth) {
        int id = (int)out->nodes.size();
        out->nodes.push_back(GqpBvhNode());
        out->max_depth = std::max(out->max_depth, depth);
        float bl[3] = {1e30f, 1e30f, 1e30f}, bh[3] = {-1e30f, -1e30f, -1e30f};
        float cl[3] = {1e30f, 1e30f, 1e30f}, ch[3] = {-1e30f, -1e30f, -1e30f};
        for (int i = lo; i < hi; ++i) {
            const float *t = fv + 9 * (size_t)order[i];
            for (int v = 0; v < 3; ++v)
                for (int k = 0; k < 3; ++k) {
                    bl[k] = std::min(bl[k], t[3 * v + k]);
                    bh[k] = std::max(bh[k], t[3 * v + k]);
                }
            for (int k = 0; k < 3; ++k) {
                cl[k] = std::min(cl[k], cen[3 * order[i] + k]);
                ch[k] = std::max(ch[k], cen[3 * order[i] + k]);
            }
        }
        GqpBvhNode n;
        for (int k = 0; k < 3; ++k) {
            n.lo[k] = bl[k] - pad - 4e-7f * std::fabs(bl[k]);
            n.hi[k] = bh[k] + pad + 4e-7f * std::fabs(bh[k]);
        }
        if (hi - lo <= leaf_size) {
            n.a = lo;
            n.b = -(hi - lo);
            out->nodes[id] = n;
            return id;
        }
        int ax = 0;
        if (ch[1] - cl[1] > ch[ax] - cl[ax]) ax = 1;
        if (ch[2] - cl[2] > ch[ax] - cl[ax]) ax = 2;
        int mid = (lo + hi) / 2;
        std::nth_element(order.begin() + lo, order.begin() + mid, order.begin() + hi,
                         [&](int a, int b) { return cen[3 * a + ax] < cen[3 * b + ax]; });
        int l = build(lo, mid, depth + 1);
        int r = build(mid, hi, depth + 1);
        n.a = l;
        n.b = r;
        out->nodes[id] = n;
        return id;
    }
};
}  // namespace

void gqp_bvh_build(const float *face_verts, int F, int leaf_size, GqpBvhHost &out) {
    out.nodes.clear();
    out.max_depth = 0;
    Builder b;
    b.fv = face_verts;
    b.leaf_size = std::max(1, leaf_size);
    b.out = &out;
    b.pad = 1e-6f;
    b.order.resize(F);
    std::iota(b.order.begin(), b.order.end(), 0);
    b.cen.resize(3 * (size_t)F);
    for (int f = 0; f < F; ++f)
        for (int k = 0; k < 3; ++k)
            b.cen[3 * f + k] = (face_verts[9 * f + k] + face_verts[9 * f + 3 + k] + face_verts[9 * f + 6 + k]) / 3.f;
    if (F > 0) b.build(0, F, 1);
    // inside a leaf keep the original face order (first-minimum rule is then free of tie handling inside a leaf)
    for (auto &n : out.nodes)
        if (n.b < 0) std::sort(b.order.begin() + n.a, b.order.begin() + n.a - n.b);
    out.tris.resize(9 * (size_t)F);
    out.face_id.resize(F);
    for (int i = 0; i < F; ++i) {
        out.face_id[i] = b.order[i];
        std::memcpy(&out.tris[9 * (size_t)i], face_verts + 9 * (size_t)b.order[i], 36);
    }
    // pre-test records (gqp_bvh.cuh: tri_lower_bound): the face's plane (unit normal through the centroid) and the
    // radius of the circle around the centroid that holds the face, both from double arithmetic, radius rounded up
    out.pre.assign(8 * (size_t)F, 0.f);
    out.extent = 0.f;
    for (int i = 0; i < F; ++i) {
        const float *t = &out.tris[9 * (size_t)i];
        D3 a = ldv(t), bb = ldv(t + 3), c = ldv(t + 6);
        D3 cen = (a + bb + c) * (1.0 / 3.0);
        D3 n = cross(bb - a, c - a);
        double l = std::sqrt(dot(n, n));
        float *q = &out.pre[8 * (size_t)i];
        float cf[3] = {(float)cen.x, (float)cen.y, (float)cen.z};
        D3 cr = {cf[0], cf[1], cf[2]};  // the ROUNDED centre is what the kernel uses: measure the radius from it
        double r = 0;
        for (D3 v : {a, bb, c}) r = std::max(r, std::sqrt(dot(v - cr, v - cr)));
        // a degenerate face gets a zero normal: the bound degrades to the sphere bound, still valid
        bool ok = l > 1e-30 && l > 1e-9 * std::max({dot(bb - a, bb - a), dot(c - a, c - a)});
        q[0] = ok ? (float)(n.x / l) : 0.f; q[1] = ok ? (float)(n.y / l) : 0.f; q[2] = ok ? (float)(n.z / l) : 0.f;
        q[3] = (float)(r * 1.0001 + 1e-7 * (std::fabs(cen.x) + std::fabs(cen.y) + std::fabs(cen.z)) + 1e-30);
        q[4] = cf[0]; q[5] = cf[1]; q[6] = cf[2]; q[7] = 0.f;
        for (int k = 0; k < 9; ++k) out.extent = std::max(out.extent, std::fabs(t[k]));
    }
}

// ---- 8-wide form of the tree (gqp_bvh.cuh: bvh_query_gang) ---------------------------------------------------------
// Collapses the binary tree: a wide node starts from a binary node's two children and keeps replacing the inner child
// with the largest box surface by that child's two children until it has 8 children or only leaves.  Boxes are the
// binary nodes' (already inflated) boxes, leaves are the binary leaves, so the set of faces and the bounds are unchanged.
namespace {
double box_area(const GqpBvhNode &n) {
    double x = (double)n.hi[0] - n.lo[0], y = (double)n.hi[1] - n.lo[1], z = (double)n.hi[2] - n.lo[2];
    return x * y + y * z + z * x;
}
// Patch record of a leaf (same layout as a face's pre-test record): unit mean normal, radius of the circle around the
// centre that holds every vertex, centre, half-thickness of the slab |n.(x - c)| <= tau that holds every vertex.  The
// faces are convex hulls of their vertices, so every point of the patch satisfies both constraints.  Normal and centre
// are rounded to fp32 FIRST; radius and thickness are measured from the rounded values in double and rounded up.
void patch_record(const GqpBvhHost &in, int first, int count, float *q) {
    D3 nsum = {0, 0, 0}, csum = {0, 0, 0};
    for (int i = first; i < first + count; ++i) {
        const float *t = &in.tris[9 * (size_t)i];
        D3 a = ldv(t), b = ldv(t + 3), c = ldv(t + 6);
        nsum = nsum + cross(b - a, c - a);
        csum = csum + a + b + c;
    }
    D3 cen = csum * (1.0 / (3.0 * count));
    double l = std::sqrt(dot(nsum, nsum));
    float nf[3] = {0.f, 0.f, 0.f};
    if (l > 1e-30) { nf[0] = (float)(nsum.x / l); nf[1] = (float)(nsum.y / l); nf[2] = (float)(nsum.z / l); }
    float cf[3] = {(float)cen.x, (float)cen.y, (float)cen.z};
    D3 nr = {nf[0], nf[1], nf[2]}, cr = {cf[0], cf[1], cf[2]};
    double r = 0, tau = 0;
    for (int i = first; i < first + count; ++i)
        for (int v = 0; v < 3; ++v) {
            D3 x = ldv(&in.tris[9 * (size_t)i + 3 * v]);
            r = std::max(r, std::sqrt(dot(x - cr, x - cr)));
            tau = std::max(tau, std::fabs(dot(nr, x - cr)));
        }
    const double pad = 1e-7 * (std::fabs(cen.x) + std::fabs(cen.y) + std::fabs(cen.z)) + 1e-30;
    q[0] = nf[0]; q[1] = nf[1]; q[2] = nf[2];
    q[3] = (float)(r * 1.0001 + pad);
    q[4] = cf[0]; q[5] = cf[1]; q[6] = cf[2];
    q[7] = (float)(tau * 1.0001 + pad);
}
int widen(const GqpBvhHost &in, int bin, int depth, GqpBvhHost &out) {
    const int id = (int)out.wide.size() / 8;
    out.wide.resize(out.wide.size() + 8, GqpBvhNode());
    out.wide_pre.resize(out.wide_pre.size() + 64, 0.f);
    out.wide_depth = std::max(out.wide_depth, depth);
    std::vector<int> kids;
    if (in.nodes[bin].b < 0) kids.push_back(bin);  // a mesh of <= leaf_size faces: the root is a leaf
    else kids = {in.nodes[bin].a, in.nodes[bin].b};
    while (kids.size() < 8) {
        int pick = -1;
        for (size_t i = 0; i < kids.size(); ++i)
            if (in.nodes[kids[i]].b >= 0 && (pick < 0 || box_area(in.nodes[kids[i]]) > box_area(in.nodes[kids[pick]])))
                pick = (int)i;
        if (pick < 0) break;
        const GqpBvhNode n = in.nodes[kids[pick]];
        kids[pick] = n.a;
        kids.push_back(n.b);
    }
    for (size_t i = 0; i < kids.size(); ++i) {
        GqpBvhNode r = in.nodes[kids[i]];
        if (r.b >= 0) {
            r.a = widen(in, kids[i], depth + 1, out);
            r.b = 1;
        } else {
            patch_record(in, r.a, -r.b, &out.wide_pre[8 * (8 * (size_t)id + i)]);
        }
        out.wide[8 * (size_t)id + i] = r;
    }
    return id;
}
}  // namespace

void gqp_bvh_widen(GqpBvhHost &bvh) {
    bvh.wide.clear();
    bvh.wide_pre.clear();
    bvh.wide_depth = 0;
    if (bvh.nodes.empty()) return;
    GqpBvhHost tmp;
    tmp.wide_depth = 0;
    widen(bvh, 0, 1, tmp);
    bvh.wide.swap(tmp.wide);
    bvh.wide_pre.swap(tmp.wide_pre);
    bvh.wide_depth = tmp.wide_depth;
}

// ---- exact culling of "outside" queries ------------------------------------------------------------------------------
// The reference's sign is sign(dot(p - c, n_f)) for the WINNING face f (.cu:236); c lies in f's plane, so it is the side
// of f's plane p is on.  For a closed, consistently outward-oriented manifold mesh and a point p OUTSIDE it, p - c lies
// in the cone spanned by the normals of the faces incident to the nearest feature; if all faces sharing a vertex have
// pairwise non-negative normal products ("nice" mesh: no crease sharper than 90 degrees, except creases that no outside
// point can be nearest to — see gqp_mesh_is_nice), that cone has a non-negative
// product with every incident normal, so the reference returns +1 for EVERY outside point, whichever tying face wins.
// Hence for a nice mesh a query can be skipped as soon as the point is known to be outside: outside the bounding box,
// or inside a grid cell that lies entirely outside the mesh.
#include <map>
#include <array>
namespace {
void nearest(const GqpBvhHost &bvh, D3 p, double &best, int &slot) {
    int stack[128], sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const GqpBvhNode &n = bvh.nodes[stack[--sp]];
        if (box_dist(n, p) > best) continue;
        if (n.b < 0) {
            for (int i = n.a; i < n.a - n.b; ++i) {
                const float *t = &bvh.tris[9 * (size_t)i];
                double d = tri_dist(p, ldv(t), ldv(t + 3), ldv(t + 6));
                if (d < best) { best = d; slot = i; }
            }
        } else {
            double dl = box_dist(bvh.nodes[n.a], p), dr = box_dist(bvh.nodes[n.b], p);
            if (dl < dr) { stack[sp++] = n.b; stack[sp++] = n.a; }
            else { stack[sp++] = n.a; stack[sp++] = n.b; }
        }
    }
}
}  // namespace

bool gqp_mesh_is_nice(const float *fv, int F) {
    if (F < 4) return false;
    std::map<std::array<float, 3>, int> vid;
    std::vector<int> idx(3 * (size_t)F);
    for (int i = 0; i < 3 * F; ++i) {
        std::array<float, 3> k = {fv[3 * i], fv[3 * i + 1], fv[3 * i + 2]};
        auto it = vid.find(k);
        if (it == vid.end()) it = vid.emplace(k, (int)vid.size()).first;
        idx[i] = it->second;
    }
    // closed + consistently oriented: every directed edge appears once and its reverse appears once
    std::map<std::pair<int, int>, int> edges, edge_face;
    for (int f = 0; f < F; ++f)
        for (int k = 0; k < 3; ++k) {
            int u = idx[3 * f + k], v = idx[3 * f + (k + 1) % 3];
            if (u == v) return false;
            if (++edges[{u, v}] > 1) return false;
            edge_face[{u, v}] = f;
        }
    for (auto &e : edges)
        if (!edges.count({e.first.second, e.first.first})) return false;
    // unit normals; no slivers
    std::vector<D3> nh(F);
    for (int f = 0; f < F; ++f) {
        D3 a = ldv(fv + 9 * f), b = ldv(fv + 9 * f + 3), c = ldv(fv + 9 * f + 6);
        D3 n = cross(b - a, c - a);
        double l = std::sqrt(dot(n, n));
        double e = std::max({dot(b - a, b - a), dot(c - a, c - a), dot(c - b, c - b)});
        if (!(l > 1e-6 * e)) return false;
        nh[f] = n * (1.0 / l);
    }
    std::vector<std::vector<int>> inc(vid.size());
    std::vector<D3> vpos(vid.size());
    for (int f = 0; f < F; ++f)
        for (int k = 0; k < 3; ++k) {
            inc[idx[3 * f + k]].push_back(f);
            vpos[idx[3 * f + k]] = ldv(fv + 9 * f + 3 * k);
        }
    // creases sharper than 90 degrees are tolerated where no OUTSIDE point can have them as its nearest feature:
    // (a) an edge whose two faces have opposed normals must be concave (the far corner of each face lies above the
    //     other's plane): the points nearest to a concave edge are behind both faces, and for a point in the valley the
    //     nearer face beats the edge by a factor <= sin(45 deg), far beyond fp32 ties;
    // (b) a vertex with an opposed pair among its faces must have an empty normal cone with a margin: for EVERY direction
    //     d some incident edge direction u has d.u >= rho, so the edge (v,u) is nearer than v by a relative
    //     rho*|u-v|/|p-v| in squared distance, which must stay above fp32 rounding (4e-6 at 2 m) so that no incident
    //     face that files p under its corner v (distance |p-v|, any face order) can win.
    for (int f = 0; f < F; ++f)
        for (int k = 0; k < 3; ++k) {
            int u = idx[3 * f + k], v = idx[3 * f + (k + 1) % 3];
            int g = edge_face[{v, u}];
            if (dot(nh[f], nh[g]) >= -1e-6) continue;
            int og = -1;  // corner of g opposite the shared edge
            for (int q = 0; q < 3; ++q)
                if (idx[3 * g + q] != u && idx[3 * g + q] != v) og = idx[3 * g + q];
            if (og < 0 || !(dot(nh[f], vpos[og] - vpos[u]) > 1e-6)) return false;  // convex (or flat-folded) sharp crease
        }
    for (size_t v = 0; v < inc.size(); ++v) {
        const std::vector<int> &fs = inc[v];
        bool sharp = false;
        for (size_t i = 0; i < fs.size() && !sharp; ++i)
            for (size_t j = i + 1; j < fs.size(); ++j)
                if (dot(nh[fs[i]], nh[fs[j]]) < -1e-6) { sharp = true; break; }
        if (!sharp) continue;
        std::vector<D3> dir;  // unit directions of the incident edges
        double emin = 1e30;
        for (int f : fs)
            for (int k = 0; k < 3; ++k) {
                int w = idx[3 * f + k];
                if (w == (int)v) continue;
                D3 e = vpos[w] - vpos[v];
                double l = std::sqrt(dot(e, e));
                emin = std::min(emin, l);
                dir.push_back(e * (1.0 / l));
            }
        // rho = radius of the largest ball about the origin inside conv{dir}: min over the hull's facets
        double rho = 1e30;
        bool any = false;
        const int n = (int)dir.size();
        for (int i = 0; i < n; ++i)
            for (int j = i + 1; j < n; ++j)
                for (int k = j + 1; k < n; ++k) {
                    D3 m = cross(dir[j] - dir[i], dir[k] - dir[i]);
                    double l = std::sqrt(dot(m, m));
                    if (l < 1e-9) continue;
                    m = m * (1.0 / l);
                    double h = dot(m, dir[i]);
                    bool below = true, above = true;
                    for (int t = 0; t < n; ++t) {
                        double sd = dot(m, dir[t]) - h;
                        if (sd > 1e-9) below = false;
                        if (sd < -1e-9) above = false;
                    }
                    if (below == above) continue;  // not a supporting plane (or everything coplanar)
                    any = true;
                    rho = std::min(rho, below ? h : -h);
                }
        if (!any || !(rho > 0.05) || !(rho * emin > 8e-6)) return false;
    }
    // outward orientation: signed volume > 0
    double vol = 0;
    for (int f = 0; f < F; ++f) vol += dot(ldv(fv + 9 * f), cross(ldv(fv + 9 * f + 3), ldv(fv + 9 * f + 6)));
    return vol > 0;
}

// ---- fp32 ties at right-angle creases -------------------------------------------------------------------------------
// The argument above is exact-arithmetic.  The reference compares SQUARED distances in fp32, so a face A ties with the
// truly nearest face B whenever d_A^2 - d_B^2 = s^2 < ~eps * d^2, i.e. when the foot point of p on B lies within
// s < sqrt(k eps) d of the feature A touches; the first-minimum rule then lets A win if its index is lower, and the
// sign is the side of A's plane.  p - c_A = d n_B + O(s), so p is BEHIND A's plane only if n_A.n_B <~ s/d ~ 1e-3: faces
// that meet at a right angle (box-like link meshes: the Allegro palm's flat top, the Shadow Hand's box links, the Panda
// fingers).  There the reference returns -1 — a "penetration" as deep as the point is far — for outside points in a
// slab of relative thickness sqrt(k eps) just behind the extended plane of A (found by the direct culling-proof test:
// Allegro palm, p 1.28 m away and 44 um under the plane of the top face; 4.5e-6 of uniformly drawn points within
// 0.25 m).  To reproduce the reference there, culling must not skip such points: this returns the planes of all faces
// that share a vertex with a face at >= 90 degrees - 0.11 degrees; the kernels keep every (sample, link) pair whose
// sample lies in  -GQP_TIE_NEG * D <= n.q - a <= GQP_TIE_POS * D  of one of them (D = an upper bound of the distance).
// Measured on the Allegro palm (33k events in 36M slab points): depth / distance <= 2.43e-4 = sqrt(2^-24), a sharp edge;
// the spacing of fp32 allows sqrt(2^-23) = 3.45e-4; GQP_TIE_NEG = 5e-4.
bool gqp_mesh_tie_planes(const float *fv, int F, int max_planes, std::vector<float> &planes,
                         std::vector<float> &anchors, int max_anchors) {
    planes.clear();
    anchors.clear();
    std::map<std::array<float, 3>, std::vector<int>> inc;
    std::vector<D3> nh(F);
    for (int f = 0; f < F; ++f) {
        D3 a = ldv(fv + 9 * f), b = ldv(fv + 9 * f + 3), c = ldv(fv + 9 * f + 6);
        D3 n = cross(b - a, c - a);
        double l = std::sqrt(dot(n, n));
        nh[f] = l > 0 ? n * (1.0 / l) : D3{0, 0, 0};
        for (int k = 0; k < 3; ++k) inc[{fv[9 * f + 3 * k], fv[9 * f + 3 * k + 1], fv[9 * f + 3 * k + 2]}].push_back(f);
    }
    std::vector<char> sus(F, 0);
    for (auto &kv : inc)
        for (int fa : kv.second)
            for (int fb : kv.second)
                if (fa != fb && dot(nh[fa], nh[fb]) <= 2e-3) sus[fa] = 1;
    for (int f = 0; f < F; ++f) {
        if (!sus[f]) continue;
        const double a0 = dot(nh[f], ldv(fv + 9 * f));
        bool dup = false;
        for (size_t i = 0; i < planes.size() && !dup; i += 4) {
            D3 m = {planes[i], planes[i + 1], planes[i + 2]};
            dup = dot(m, nh[f]) > 1.0 - 1e-9 && std::fabs(planes[i + 3] - a0) < 1e-7;
        }
        if (dup) continue;
        planes.push_back((float)nh[f].x); planes.push_back((float)nh[f].y); planes.push_back((float)nh[f].z);
        planes.push_back((float)a0);
        if ((int)planes.size() / 4 > max_planes) return false;
    }
    // anchors: vertices of the tie faces.  A tie needs the nearest feature on the boundary of a tie face, so the nearest
    // of these vertices bounds the distance of a slab point from above far better than the link's bounding sphere does.
    // More than max_anchors: keep every k-th (any subset of mesh vertices gives a valid upper bound).
    std::vector<std::array<float, 3>> av;
    for (int f = 0; f < F; ++f) {
        if (!sus[f]) continue;
        for (int k = 0; k < 3; ++k) {
            std::array<float, 3> v = {fv[9 * f + 3 * k], fv[9 * f + 3 * k + 1], fv[9 * f + 3 * k + 2]};
            if (std::find(av.begin(), av.end(), v) == av.end()) av.push_back(v);
        }
    }
    const size_t stride = av.size() > (size_t)max_anchors ? (av.size() + max_anchors - 1) / max_anchors : 1;
    for (size_t i = 0; i < av.size(); i += stride) {
        anchors.push_back(av[i][0]); anchors.push_back(av[i][1]); anchors.push_back(av[i][2]); anchors.push_back(0.f);
    }
    return true;
}

// bit = 1 when the whole cell is outside the (nice) mesh with a 10 micron margin.  Cell (x,y,z) -> bit
// ((z*dim[1]+y)*dim[0]+x).
// A cell is FREE when the surface stays farther from its centre than its circumradius (+ margin).  A free cell is convex
// and untouched by the surface, so it lies in ONE region of space, and two free cells sharing a face lie in the same
// one.  Free cells of the outermost layer reach beyond the mesh's bounding box (the grid overhangs it by at least half
// a cell on every side), hence are outside; outside-ness is flood-filled from them through face-adjacent free cells.
// (Judging a cell by the side of its centre's nearest face alone is wrong for meshes with nested shells, e.g. the
// Shadow Hand's thumb base: between two outward-oriented shells the nearest face may be the inner one.)
void gqp_outside_grid_build(const GqpBvhHost &bvh, const float origin[3], float cell, const int dim[3],
                            std::vector<uint32_t> &bits) {
    size_t ncell = (size_t)dim[0] * dim[1] * dim[2];
    bits.assign((ncell + 31) / 32, 0u);
    if (bvh.nodes.empty()) return;
    const double r = 0.5 * std::sqrt(3.0) * cell * 1.001 + 1e-5;  // circumradius + slack for the fp32 cell index + margin
    std::vector<unsigned char> state(ncell, 0);                   // 0 = touched by the surface, 1 = free, 2 = outside
    for (int z = 0; z < dim[2]; ++z)
        for (int y = 0; y < dim[1]; ++y)
            for (int x = 0; x < dim[0]; ++x) {
                D3 c = {origin[0] + (x + 0.5) * (double)cell, origin[1] + (y + 0.5) * (double)cell,
                        origin[2] + (z + 0.5) * (double)cell};
                double best = 1e30;
                int slot = -1;
                nearest(bvh, c, best, slot);
                if (slot < 0 || best <= r) continue;
                const float *t = &bvh.tris[9 * (size_t)slot];
                D3 a = ldv(t), b = ldv(t + 3), cc = ldv(t + 6);
                if (dot(c - a, cross(b - a, cc - a)) <= 0) continue;  // centre is inside
                state[((size_t)z * dim[1] + y) * dim[0] + x] = 1;
            }
    std::vector<size_t> todo;
    auto visit = [&](int x, int y, int z) {
        if (x < 0 || y < 0 || z < 0 || x >= dim[0] || y >= dim[1] || z >= dim[2]) return;
        size_t i = ((size_t)z * dim[1] + y) * dim[0] + x;
        if (state[i] != 1) return;
        state[i] = 2;
        todo.push_back(i);
    };
    for (int z = 0; z < dim[2]; ++z)
        for (int y = 0; y < dim[1]; ++y)
            for (int x = 0; x < dim[0]; ++x)
                if (x == 0 || y == 0 || z == 0 || x == dim[0] - 1 || y == dim[1] - 1 || z == dim[2] - 1) visit(x, y, z);
    while (!todo.empty()) {
        size_t i = todo.back();
        todo.pop_back();
        int x = (int)(i % dim[0]), y = (int)((i / dim[0]) % dim[1]), z = (int)(i / ((size_t)dim[0] * dim[1]));
        visit(x - 1, y, z); visit(x + 1, y, z); visit(x, y - 1, z); visit(x, y + 1, z); visit(x, y, z - 1); visit(x, y, z + 1);
    }
    for (size_t i = 0; i < ncell; ++i)
        if (state[i] == 2) bits[i >> 5] |= 1u << (i & 31);
}
