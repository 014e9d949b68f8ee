// Meshes with a BVH: the exact-winner accelerated form of K1, and O1 = ObjectModel.cal_distance
// (core/object_model.py:186-255, TORCHSDF branch) for all objects of a device in one launch.
// A gang of 8 lanes owns a query point and walks the 8-wide BVH (gqp_bvh.cuh: bvh_query_gang; child records 32 B, read
// as two float4 through L1/L2; a 50k-face object mesh is 1.8 MB of faces + 1.6 MB of pre-test records + 0.4 MB of nodes,
// L2 resident).  Arithmetic per (point, face) is the non-contracted reference order,
// so distances, signs and closest points are bit-identical to gqp_sdf_forward and to the host-compiled reference.
#include "gqp_bvh.cuh"
#include "graspqp_b200.h"

struct GqpObjSet {
    int n_obj = 0;
    std::vector<const GqpMesh *> meshes;
    const GqpBvhNode **d_nodes = nullptr;
    const float **d_tris = nullptr;
    const int32_t **d_face_id = nullptr;
    const float **d_pre = nullptr;
    const float **d_wpre = nullptr;
    float *d_extent = nullptr;
};

#define MESH_WARPS 4
#ifndef MESH_MIN_BLOCKS
#define MESH_MIN_BLOCKS 8
#endif
#define MESH_GANGS (MESH_WARPS * 32 / GQP_GANG)
// queries per warp: enough for the 4 gangs of a warp to even out searches of different length
static int mesh_qpw(long long n_queries) {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("GQP_MESH_QPW");
        v = e ? atoi(e) : 0;
    }
    if (v >= 4) return v;
    return n_queries >= 200000 ? 16 : 8;  // measured on B200: 98k queries prefer 8, 393k prefer 16
}

__global__ void __launch_bounds__(MESH_WARPS * 32, MESH_MIN_BLOCKS)
mesh_sdf_kernel(const GqpBvhNode *__restrict__ nodes, const float *__restrict__ tris,
                const int32_t *__restrict__ face_id, const float *__restrict__ pre,
                const float *__restrict__ wpre, float extent, const float *__restrict__ points, int P, int qpw, float *__restrict__ dist2,
                int32_t *__restrict__ sign, float *__restrict__ normals, float *__restrict__ clst,
                int32_t *__restrict__ face_out) {
    __shared__ int2 stk[MESH_GANGS][GQP_GANG_STACK];
    const long long w = (long long)blockIdx.x * MESH_WARPS + (threadIdx.x >> 5);
    const int q0 = (int)min((long long)P, w * qpw), q1 = (int)min((long long)P, (w + 1) * qpw);
    bvh_query_gang<true>(
        q0, q1, stk[threadIdx.x / GQP_GANG],
        [&](int i, V3 &p, BvhTree &tr) {
            p = ld3(points + 3 * (size_t)i);
            tr = BvhTree{nodes, wpre, tris, face_id, pre, extent};
        },
        [&](int i, V3 &p, BvhTree &, const SdfBest &best) {
            if ((threadIdx.x & (GQP_GANG - 1)) == 0) {
                int sg;
                V3 un;
                sdf_finish<true>(p, best, sg, un);
                dist2[i] = best.d2;
                sign[i] = sg;
                st3(normals + 3 * (size_t)i, un);
                st3(clst + 3 * (size_t)i, best.c);
                if (face_out) face_out[i] = best.face;
            }
            return false;
        });
}

// O1 forward: query i belongs to object i / (batch_each * n_pts)
__global__ void __launch_bounds__(MESH_WARPS * 32, MESH_MIN_BLOCKS)
object_distance_kernel(const GqpBvhNode *const *__restrict__ nodes, const float *const *__restrict__ tris,
                       const int32_t *const *__restrict__ face_id, const float *const *__restrict__ pre,
                       const float *const *__restrict__ wpre, const float *__restrict__ extent, int per_obj, int n, int qpw, const float *__restrict__ x,
                       float scale, float *__restrict__ distance, float *__restrict__ normals,
                       float *__restrict__ clst, int32_t *__restrict__ sign) {
    __shared__ int2 stk[MESH_GANGS][GQP_GANG_STACK];
    const long long w = (long long)blockIdx.x * MESH_WARPS + (threadIdx.x >> 5);
    const int q0 = (int)min((long long)n, w * qpw), q1 = (int)min((long long)n, (w + 1) * qpw);
    bvh_query_gang<true>(
        q0, q1, stk[threadIdx.x / GQP_GANG],
        [&](int i, V3 &p, BvhTree &tr) {
            const int o = i / per_obj;
            p = ld3(x + 3 * (size_t)i);
            p = V3{p.x / scale, p.y / scale, p.z / scale};  // object_model.py:215-216
            tr = BvhTree{nodes[o], wpre[o], tris[o], face_id[o], pre[o], extent[o]};
        },
        [&](int i, V3 &p, BvhTree &, const SdfBest &best) {
            if ((threadIdx.x & (GQP_GANG - 1)) == 0) {
                int sg;
                V3 un;
                sdf_finish<true>(p, best, sg, un);
                float d = sqrtf(best.d2 + 1e-8f) * (float)(-sg);  // :243
                distance[i] = d * scale;                          // :249
                st3(normals + 3 * (size_t)i, un * (float)sg);     // :246
                st3(clst + 3 * (size_t)i, best.c);
                sign[i] = sg;
            }
            return false;
        });
}

// O1 backward: TorchSDF backward (2 g' (p-c), sdf.py:60-66) chained through sqrt(d2+1e-8)*(-sign)*scale and p = x/scale
__global__ void object_distance_backward_kernel(const float *__restrict__ grad_distance, const float *__restrict__ x,
                                                const float *__restrict__ clst, const int32_t *__restrict__ sign,
                                                int n, float scale, float *__restrict__ grad_x) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 p = ld3(x + 3 * (size_t)i);
    p = V3{p.x / scale, p.y / scale, p.z / scale};
    V3 dv = p - ld3(clst + 3 * (size_t)i);
    float d2 = dot3(dv, dv);
    float g = grad_distance[i] * (float)(-sign[i]) / sqrtf(d2 + 1e-8f);
    st3(grad_x + 3 * (size_t)i, dv * g);
}

extern "C" int gqp_mesh_create(const float *face_verts_host, int F, GqpMesh **out) {
    if (!face_verts_host || F <= 0 || !out) return gqp_set_error(GQP_ERR_ARG, "gqp_mesh_create: need F > 0");
    GqpBvhHost bvh;
    gqp_bvh_build(face_verts_host, F, GQP_GANG, bvh);
    gqp_bvh_widen(bvh);
    if (7 * bvh.wide_depth + 1 > GQP_GANG_STACK)
        return gqp_set_error(GQP_ERR_LIMIT, "gqp_mesh_create: BVH depth %d exceeds the traversal stack", bvh.wide_depth);
    GqpMesh *m = new GqpMesh();
    m->F = F;
    m->n_nodes = (int)bvh.wide.size();
    m->max_depth = bvh.wide_depth;
    m->d_nodes = (GqpBvhNode *)gqp_dev_alloc_upload(bvh.wide.data(), bvh.wide.size() * sizeof(GqpBvhNode));
    m->d_tris = (float *)gqp_dev_alloc_upload(bvh.tris.data(), bvh.tris.size() * sizeof(float));
    m->d_face_id = (int32_t *)gqp_dev_alloc_upload(bvh.face_id.data(), bvh.face_id.size() * sizeof(int32_t));
    m->d_pre = (float *)gqp_dev_alloc_upload(bvh.pre.data(), bvh.pre.size() * sizeof(float));
    m->d_wpre = (float *)gqp_dev_alloc_upload(bvh.wide_pre.data(), bvh.wide_pre.size() * sizeof(float));
    m->extent = bvh.extent;
    if (!m->d_nodes || !m->d_tris || !m->d_face_id || !m->d_pre || !m->d_wpre) {
        gqp_mesh_destroy(m);
        return gqp_set_error(GQP_ERR_CUDA, "gqp_mesh_create: device allocation failed");
    }
    *out = m;
    return GQP_OK;
}
extern "C" void gqp_mesh_destroy(GqpMesh *m) {
    if (!m) return;
    gqp_dev_free(m->d_nodes); gqp_dev_free(m->d_tris); gqp_dev_free(m->d_face_id); gqp_dev_free(m->d_pre);
    gqp_dev_free(m->d_wpre);
    delete m;
}
extern "C" int gqp_mesh_num_faces(const GqpMesh *m) { return m ? m->F : 0; }

extern "C" int gqp_mesh_sdf_forward(const GqpMesh *m, const float *points, int P, float *dist2, int32_t *sign,
                                    float *normals, float *clst, int32_t *face, void *stream) {
    if (!m || P < 0) return gqp_set_error(GQP_ERR_ARG, "gqp_mesh_sdf_forward: bad argument");
    if (P == 0) return GQP_OK;
    if (!points || !dist2 || !sign || !normals || !clst) return gqp_set_error(GQP_ERR_ARG, "gqp_mesh_sdf_forward: null pointer");
    const int qpw = mesh_qpw(P);
    GQP_LAUNCH(mesh_sdf_kernel, (P + MESH_WARPS * qpw - 1) / (MESH_WARPS * qpw), MESH_WARPS * 32, 0,
               (cudaStream_t)stream, m->d_nodes, m->d_tris, m->d_face_id, m->d_pre, m->d_wpre, m->extent, points, P, qpw,
               dist2, sign, normals, clst, face);
    GQP_CHECK_LAUNCH("gqp_mesh_sdf_forward");
    return GQP_OK;
}

extern "C" int gqp_objset_create(const GqpMesh *const *meshes_host, int n_obj, GqpObjSet **out) {
    if (!meshes_host || n_obj <= 0 || !out) return gqp_set_error(GQP_ERR_ARG, "gqp_objset_create: bad argument");
    GqpObjSet *s = new GqpObjSet();
    s->n_obj = n_obj;
    std::vector<const GqpBvhNode *> n(n_obj);
    std::vector<const float *> t(n_obj);
    std::vector<const int32_t *> f(n_obj);
    std::vector<const float *> pr(n_obj), wp(n_obj);
    std::vector<float> ex(n_obj);
    for (int o = 0; o < n_obj; ++o) {
        if (!meshes_host[o]) {
            delete s;
            return gqp_set_error(GQP_ERR_ARG, "gqp_objset_create: mesh %d is null", o);
        }
        s->meshes.push_back(meshes_host[o]);
        n[o] = meshes_host[o]->d_nodes; t[o] = meshes_host[o]->d_tris; f[o] = meshes_host[o]->d_face_id;
        pr[o] = meshes_host[o]->d_pre; wp[o] = meshes_host[o]->d_wpre; ex[o] = meshes_host[o]->extent;
    }
    s->d_nodes = (const GqpBvhNode **)gqp_dev_alloc_upload(n.data(), n_obj * sizeof(void *));
    s->d_tris = (const float **)gqp_dev_alloc_upload(t.data(), n_obj * sizeof(void *));
    s->d_face_id = (const int32_t **)gqp_dev_alloc_upload(f.data(), n_obj * sizeof(void *));
    s->d_pre = (const float **)gqp_dev_alloc_upload(pr.data(), n_obj * sizeof(void *));
    s->d_wpre = (const float **)gqp_dev_alloc_upload(wp.data(), n_obj * sizeof(void *));
    s->d_extent = (float *)gqp_dev_alloc_upload(ex.data(), n_obj * sizeof(float));
    if (!s->d_nodes || !s->d_tris || !s->d_face_id || !s->d_pre || !s->d_wpre || !s->d_extent) {
        gqp_objset_destroy(s);
        return gqp_set_error(GQP_ERR_CUDA, "gqp_objset_create: device allocation failed");
    }
    *out = s;
    return GQP_OK;
}
extern "C" void gqp_objset_destroy(GqpObjSet *s) {
    if (!s) return;
    gqp_dev_free((void *)s->d_nodes); gqp_dev_free((void *)s->d_tris); gqp_dev_free((void *)s->d_face_id);
    gqp_dev_free((void *)s->d_pre); gqp_dev_free((void *)s->d_wpre); gqp_dev_free((void *)s->d_extent);
    delete s;
}

extern "C" int gqp_object_distance_forward(const GqpObjSet *s, int batch_each, int n_pts, const float *x, float scale,
                                           float *distance, float *normals, float *clst, int32_t *sign,
                                           void *stream) {
    if (!s || batch_each < 0 || n_pts < 0 || !(scale > 0.f))
        return gqp_set_error(GQP_ERR_ARG, "gqp_object_distance_forward: bad argument");
    long long n = (long long)s->n_obj * batch_each * n_pts;
    if (n == 0) return GQP_OK;
    if (n > 0x7fffffffLL) return gqp_set_error(GQP_ERR_LIMIT, "gqp_object_distance_forward: too many queries");
    if (!x || !distance || !normals || !clst || !sign)
        return gqp_set_error(GQP_ERR_ARG, "gqp_object_distance_forward: null pointer");
    const int qpw = mesh_qpw(n);
    GQP_LAUNCH(object_distance_kernel, (int)((n + MESH_WARPS * qpw - 1) / (MESH_WARPS * qpw)), MESH_WARPS * 32, 0,
               (cudaStream_t)stream, s->d_nodes, s->d_tris, s->d_face_id, s->d_pre, s->d_wpre, s->d_extent, batch_each * n_pts,
               (int)n, qpw, x, scale, distance, normals, clst, sign);
    GQP_CHECK_LAUNCH("gqp_object_distance_forward");
    return GQP_OK;
}

extern "C" int gqp_object_distance_backward(const float *grad_distance, const float *x, const float *clst,
                                            const int32_t *sign, int n, float scale, float *grad_x, void *stream) {
    if (n < 0 || !(scale > 0.f)) return gqp_set_error(GQP_ERR_ARG, "gqp_object_distance_backward: bad argument");
    if (n == 0) return GQP_OK;
    if (!grad_distance || !x || !clst || !sign || !grad_x)
        return gqp_set_error(GQP_ERR_ARG, "gqp_object_distance_backward: null pointer");
    GQP_LAUNCH(object_distance_backward_kernel, (n + 255) / 256, 256, 0, (cudaStream_t)stream, grad_distance, x, clst,
               sign, n, scale, grad_x);
    GQP_CHECK_LAUNCH("gqp_object_distance_backward");
    return GQP_OK;
}
