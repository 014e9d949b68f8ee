// Internal layout of the opaque handles of include/graspqp_b200.h.
#pragma once
#include <stdint.h>

#include <vector>

#define GQP_MAX_FRAMES 48
#define GQP_MAX_LINKS 32
#define GQP_MAX_DOFS 32
#define GQP_MAX_CHILDREN 8
#define GQP_MAX_MLP_HIDDEN 16
#define GQP_MAX_MLP_OUT 7

// ---- BVH over a triangle mesh (built on the host, gqp_bvh.cpp) ----------------------------------------------------
// Node = 32 bytes: box min (3 floats), `a`, box max (3 floats), `b`.
//   inner node: a = index of left child, b = index of right child (both > 0), count == 0 is encoded as b >= 0
//   leaf      : a = first slot in the reordered triangle array, b = -(count)  (count >= 1)
// Boxes are inflated by a few ulps so that a box's fp32 lower bound never exceeds the fp32 distance of a face inside.
struct GqpBvhNode {
    float lo[3];
    int32_t a;
    float hi[3];
    int32_t b;
};

struct GqpBvhHost {
    std::vector<GqpBvhNode> nodes;   // nodes[0] = root
    std::vector<float> tris;         // (F,9) vertices in leaf order
    std::vector<int32_t> face_id;    // (F) original face index of slot i
    std::vector<float> pre;          // (F,8) per slot: unit normal xyz, bounding-circle radius, centroid xyz, 0
    float extent = 0.f;              // max |coordinate| of the mesh
    int max_depth = 0;
    std::vector<GqpBvhNode> wide;    // 8-wide form (gqp_bvh_widen): 8 child records per wide node, wide node 0 = root
    std::vector<float> wide_pre;     // (8 per wide node, 8) patch record of each leaf child (gqp_bvh.cpp: patch_record)
    int wide_depth = 0;
};
// leaf_size: max faces per leaf
void gqp_bvh_build(const float *face_verts, int F, int leaf_size, GqpBvhHost &out);
// fills bvh.wide / bvh.wide_depth from bvh.nodes (leaf_size must be <= 8)
void gqp_bvh_widen(GqpBvhHost &bvh);

// Exact culling support (gqp_bvh.cpp): a "nice" mesh is a closed, outward-oriented manifold without creases sharper
// than 90 degrees; the reference's sign heuristic returns +1 for every point outside such a mesh.
bool gqp_mesh_is_nice(const float *face_verts, int F);
// Planes under which the "nice mesh => +1 outside" argument fails IN FP32 (gqp_bvh.cpp: gqp_mesh_tie_planes):
// 4 floats per plane (unit normal, offset n.a); false when the mesh has more than max_planes of them.
bool gqp_mesh_tie_planes(const float *face_verts, int F, int max_planes, std::vector<float> &planes,
                         std::vector<float> &anchors, int max_anchors);
void gqp_outside_grid_build(const GqpBvhHost &bvh, const float origin[3], float cell, const int dim[3],
                            std::vector<uint32_t> &bits);

struct GqpMesh {
    int F = 0;
    int n_nodes = 0;
    int max_depth = 0;
    GqpBvhNode *d_nodes = nullptr;  // device: wide records, 8 per wide node
    float *d_tris = nullptr;        // device (F,9) leaf order
    int32_t *d_face_id = nullptr;   // device (F)
    float *d_pre = nullptr;         // device (F,8) pre-test records in leaf order
    float *d_wpre = nullptr;        // device (8 per wide node, 8) patch records of leaf children
    float extent = 0.f;
};

// Device-side view of the hand tables (one contiguous blob, see gqp_hand.cu)
struct GqpHandDev {
    int32_t n_frames, n_dofs, n_links, n_cand, n_spheres;
    int32_t frame_parent[GQP_MAX_FRAMES], frame_jtype[GQP_MAX_FRAMES], frame_dof[GQP_MAX_FRAMES];
    int32_t frame_link[GQP_MAX_FRAMES];  // mesh-link id hanging on the frame, or -1
    float frame_axis[GQP_MAX_FRAMES][3];
    float frame_gain[GQP_MAX_FRAMES];    // joint value = gain * parameter (1 unless GqpHandDesc::frame_gain_host says otherwise)
    float frame_joff[GQP_MAX_FRAMES][12];
    float frame_loff[GQP_MAX_FRAMES][12];
    int32_t link_frame[GQP_MAX_LINKS];
    // tree shape for the warp-per-candidate kernels (gqp_fkw.cu): depth of every frame (roots 0), its children in
    // frame order, the deepest level
    int32_t frame_depth[GQP_MAX_FRAMES], frame_nchild[GQP_MAX_FRAMES], frame_child[GQP_MAX_FRAMES][GQP_MAX_CHILDREN];
    int32_t max_depth, max_children;
    float joints_lower[GQP_MAX_DOFS], joints_upper[GQP_MAX_DOFS];
    // joint-coupling network (GqpHandDesc::mlp_*); mlp_hidden == 0: frame_dof indexes the pose row directly
    int32_t mlp_hidden, mlp_out;
    float mlp_lo, mlp_hi;
    float mlp_w1[GQP_MAX_MLP_HIDDEN], mlp_b1[GQP_MAX_MLP_HIDDEN];
    float mlp_w2[GQP_MAX_MLP_OUT][GQP_MAX_MLP_HIDDEN], mlp_b2[GQP_MAX_MLP_OUT];
};

struct GqpPenTables;  // gqp_pen.cu

struct GqpHand {
    GqpHandDev host;              // host copy
    GqpHandDev *d_dev = nullptr;  // device copy
    float *d_cand = nullptr, *d_cand_normal = nullptr;  // (n_cand,3)
    int32_t *d_cand_link = nullptr;                      // (n_cand)
    float *d_spheres = nullptr;                          // (n_spheres,4)
    int32_t *d_sphere_link = nullptr;                    // (n_spheres)
    std::vector<int32_t> face_offset;                    // host (n_links+1)
    GqpPenTables *pen = nullptr;                         // per-link BVHs + sign grids (device blob)
};

int gqp_pen_tables_create(GqpHand *h, const float *face_verts_host);
void gqp_pen_tables_destroy(GqpHand *h);
