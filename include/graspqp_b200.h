/* graspqp_b200 — C ABI of the B200-native GraspQP hot path (libgraspqp_b200.so).
 *
 * The reference (leggedrobotics/graspqp) has no plugin ABI: its seam is Python duck typing plus ONE pybind module,
 * torchsdf._C (thirdparty/TorchSDF/torchsdf/csrc/bindings.cpp:22-27).  This header is what a binding for the hot path
 * `per-iteration grasp-energy evaluation + MALA* update` binds instead; each entry point cites what it replaces.
 *
 * Conventions
 *   - every pointer argument is a DEVICE pointer owned by the caller unless its name ends in `_host`;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - no allocation happens after a *_create call returns; handles are immutable device-resident tables;
 *   - return value 0 = ok, negative = error; gqp_last_error() gives the message of the calling thread's last error;
 *   - no exceptions cross the ABI; one host thread per device (as in the reference: single Python thread).
 *   - all floating point is fp32 (the QP's 6x6 Schur solve uses fp64 internally), indices are int32 on the device.
 */
#ifndef GRASPQP_B200_H
#define GRASPQP_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GQP_OK 0
#define GQP_ERR_ARG (-1)
#define GQP_ERR_CUDA (-2)
#define GQP_ERR_LIMIT (-3)

typedef struct GqpMesh GqpMesh; /* triangle mesh + BVH + far-field sign grid, device resident */
typedef struct GqpObjSet GqpObjSet; /* the objects of one device: table of mesh handles for single-launch queries */
typedef struct GqpHand GqpHand; /* kinematic tree, contact candidates, penetration spheres, link meshes */

int gqp_abi_version(void);
const char *gqp_last_error(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches counter) */
uint64_t gqp_launch_count(void);

/* ---- K1: TorchSDF contract --------------------------------------------------------------------------------------
 * replaces torchsdf._C.unbatched_triangle_distance_forward_cuda / _backward_cuda
 *   (thirdparty/TorchSDF/torchsdf/csrc/bindings.cpp:22-27, unbatched_triangle_distance_cuda.cu:176-270).
 * points (P,3), face_verts (F,3,3) contiguous.  Outputs dist2 (P), sign (P) int32 in {+1,-1}, normals (P,3) =
 * unit(p - closest), clst (P,3).  Winner = first minimum in face order; arithmetic is not FMA-contracted so results
 * are bit-identical to the host-compiled reference kernel.  Brute force over face tiles staged in shared memory. */
int gqp_sdf_forward(const float *points, const float *face_verts, int P, int F, float *dist2, int32_t *sign,
                    float *normals, float *clst, void *stream);
int gqp_sdf_backward(const float *grad_dist, const float *points, const float *clst, int P, float *grad_points,
                     void *stream);

/* ---- meshes with an exact acceleration structure ------------------------------------------------------------------
 * gqp_mesh_create builds (on the host, once) a BVH over the faces and uploads it; gqp_mesh_sdf_forward answers the
 * same query as gqp_sdf_forward with the SAME winner (lexicographic (dist2, face index) minimum), visiting only the
 * nodes whose box can hold a closer face.  face (P) int32 receives the winning face index (may be NULL).
 * replaces the `face_verts`/`wp.Mesh` members of ObjectModel / HandModel (object_model.py:147-161,
 * hand_model.py:351-372). */
int gqp_mesh_create(const float *face_verts_host, int F, GqpMesh **out);
void gqp_mesh_destroy(GqpMesh *m);
int gqp_mesh_num_faces(const GqpMesh *m);
int gqp_mesh_sdf_forward(const GqpMesh *m, const float *points, int P, float *dist2, int32_t *sign, float *normals,
                         float *clst, int32_t *face, void *stream);

/* ---- O1: ObjectModel.cal_distance (object_model.py:186-255, TORCHSDF branch) -----------------------------------------
 * An object set is the list of meshes of the objects a device optimises (ObjectModel.initialize, object_model.py:70-184);
 * the meshes must outlive the set.  x (n_obj*batch_each, n_pts, 3); object o owns rows [o*batch_each, (o+1)*batch_each).
 * distance = sqrt(d2+1e-8)*(-sign)*scale, normals = unit(p-c)*sign (outward object normal); clst and sign are kept for
 * the backward:  grad_x = grad_distance * (-sign) * (p-c)/sqrt(d2+1e-8)  (sdf.py:60-66 chained with :243-249).
 * All objects are answered by ONE launch (the reference launches TorchSDF once per object, :217-220). */
int gqp_objset_create(const GqpMesh *const *meshes_host, int n_obj, GqpObjSet **out);
void gqp_objset_destroy(GqpObjSet *s);
int gqp_object_distance_forward(const GqpObjSet *s, int batch_each, int n_pts, const float *x, float scale,
                                float *distance, float *normals, float *clst, int32_t *sign, void *stream);
int gqp_object_distance_backward(const float *grad_distance, const float *x, const float *clst, const int32_t *sign,
                                 int n, float scale, float *grad_x, void *stream);

/* ---- hand description --------------------------------------------------------------------------------------------
 * All arrays are HOST pointers, copied.  Frames are in the DFS order of pytorch_kinematics' chain
 * (thirdparty/pytorch_kinematics/pytorch_kinematics/chain.py:128-138), parent index < own index.
 * jtype: 0 fixed, 1 revolute, 2 prismatic.  frame_dof: column of hand_pose[:,9:] driving the joint, or -1.
 * joff / loff: joint.offset and link.offset as row-major 3x4 (frame.py:98-114).
 * link l (a "mesh link", hand_model.py:219-246) hangs on frame link_frame[l]; its faces are
 * face_verts[face_offset[l] .. face_offset[l+1]).  cand*: contact candidates in link frames with their link id,
 * in the reference's concatenation order (hand_model.py:1220-1267).  spheres: xyz + radius per penetration keypoint,
 * sphere_link ascending (hand_model.py:308-324, 989-1040). */
typedef struct GqpHandDesc {
    int32_t n_frames, n_dofs, n_links, n_cand, n_spheres;
    const int32_t *frame_parent_host, *frame_jtype_host, *frame_dof_host;
    const float *frame_axis_host; /* (n_frames,3) */
    const float *frame_joff_host; /* (n_frames,3,4) */
    const float *frame_loff_host; /* (n_frames,3,4) */
    const int32_t *link_frame_host;  /* (n_links) */
    const int32_t *face_offset_host; /* (n_links+1) */
    const float *face_verts_host;    /* (face_offset[n_links],3,3) */
    const float *cand_host, *cand_normal_host; /* (n_cand,3) */
    const int32_t *cand_link_host;             /* (n_cand) */
    const float *spheres_host;                 /* (n_spheres,4) */
    const int32_t *sphere_link_host;           /* (n_spheres) */
    const float *joints_lower_host, *joints_upper_host; /* (n_dofs) */
    /* Joint-coupling network of the Robotiq 2F-140 (hands/robotiq2.py:9-13, utils/fk.py:23-51); mlp_hidden == 0: none.
     * With it n_dofs must be 1 and frame_dof indexes the expanded joint vector
     *   q = [c, W2 tanh(w1 c + b1) + b2],  c = clamp(pose[9], mlp_lo, mlp_hi)        (1 + mlp_out entries)
     * instead of the pose row.  w1, b1: (mlp_hidden); w2: (mlp_out, mlp_hidden) row-major; b2: (mlp_out).
     * Limits: mlp_hidden <= 16, mlp_out <= 7. */
    int32_t mlp_hidden, mlp_out;
    float mlp_lo, mlp_hi;
    const float *mlp_w1_host, *mlp_b1_host, *mlp_w2_host, *mlp_b2_host;
    /* (n_frames) or NULL = all 1: the joint of frame f takes gain * hand_pose[:, 9 + frame_dof[f]].  Mimic joints of the
     * ability hand (hands/ability_hand.py:10-29: the second joint of a finger = 1.05851325 x the first); a Schunk-style
     * mirrored finger would be gain -1 (hands/schunk.py:31-38). */
    const float *frame_gain_host;
} GqpHandDesc;

int gqp_hand_create(const GqpHandDesc *desc, GqpHand **out);
void gqp_hand_destroy(GqpHand *h);

/* ---- H1/H2: HandModel.set_parameters' kinematics (hand_model.py:856-865, chain.py:146-168, frame.py:98-114) -------
 * pose (Bt, 9+n_dofs): translation, ortho6d, joint angles.  link_tf (Bt, n_links, 12): WORLD transform of every mesh
 * link, row-major 3x4, i.e. [R_g R_l | R_g t_l + t_g].  Includes the reference's +-0.01 rad revolute perturbation.
 * gqp_fk_backward maps d(loss)/d(link_tf) to d(loss)/d(pose) (replaces autograd through the chain, fit.py:440);
 * grad_pose is OVERWRITTEN with that plus grad_pose_extra (Bt, 9+n_dofs; may be NULL, e.g. the joint-limit term). */
int gqp_fk_forward(const GqpHand *h, const float *pose, int Bt, float *link_tf, void *stream);
int gqp_fk_backward(const GqpHand *h, const float *pose, const float *grad_link_tf, const float *grad_pose_extra,
                    int Bt, float *grad_pose, void *stream);

/* ---- H3: contact gather (hand_model.py:787-831, 1220-1267) ---------------------------------------------------------
 * idx (Bt,n_c) int32 indices into the candidate table; cpts/cnrm (Bt,n_c,3) world position / normal.
 * backward ACCUMULATES into grad_link_tf (Bt,n_links,12) (caller zeroes it once per step). */
int gqp_contacts_forward(const GqpHand *h, const float *link_tf, const int32_t *idx, int Bt, int n_c, float *cpts,
                         float *cnrm, void *stream);
int gqp_contacts_backward(const GqpHand *h, const int32_t *idx, const float *grad_cpts, const float *grad_cnrm, int Bt,
                          int n_c, float *grad_link_tf, void *stream);

/* ---- H1-H3 fused, a warp per candidate (what the engine launches): gqp_fk_forward + gqp_contacts_forward in one launch,
 * and gqp_contacts_backward + gqp_fk_backward in one launch.  idx may be NULL (kinematics only).  The backward READS
 * grad_link_tf (Bt,n_links,12; what the penetration stages accumulated; may be NULL = zero), adds this candidate's
 * contact terms in contact order and OVERWRITES grad_pose with the total plus grad_pose_extra (may be NULL); every
 * reduction has a fixed order (no atomics).  Hands with more than 32 frames fall back to the four separate kernels. */
int gqp_fk_contacts_forward(const GqpHand *h, const float *pose, const int32_t *idx, int Bt, int n_c, float *link_tf,
                            float *cpts, float *cnrm, void *stream);
int gqp_fk_contacts_backward(const GqpHand *h, const float *pose, const int32_t *idx, const float *grad_cpts,
                             const float *grad_cnrm, float *grad_link_tf, const float *grad_pose_extra, int Bt, int n_c,
                             float *grad_pose, void *stream);

/* ---- H4: HandModel.cal_distance (hand_model.py:875-987) --------------------------------------------------------------
 * x (n_obj, n_pts, 3): one point set per object, shared by its batch_each candidates (the reference materialises the
 * repeat, object_model.py:181-184; x_stride_candidates != 0 selects a per-candidate (Bt,n_pts,3) set instead).
 * dist (Bt,n_pts) = max over links of sqrt(d2+1e-8)*(-sign)   [positive inside], may be NULL;
 * e_pen (Bt) = sum_p relu(dist) (energy.py:57-62), may be NULL;
 * scratch (required): workspace of gqp_hand_distance_scratch_floats(...) floats — per-chunk partial sums (deterministic
 *   two-stage sum), bounding spheres of the 32-sample chunks, per-candidate link masks;
 * grad_link_tf (Bt,n_links,12), may be NULL: ACCUMULATES w_pen * d(e_pen)/d(link_tf).
 * With dist == NULL, (point, link) pairs whose sign is provably +1 (far-field sign grid built at gqp_hand_create) are
 * skipped: they cannot contribute to relu(max).  With dist != NULL every pair is searched. */
int gqp_hand_distance(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each, int n_pts,
                      int per_candidate_points, float *dist, float *e_pen, float *scratch, float w_pen,
                      float *grad_link_tf, void *stream);

/* cal_distance for autograd callers (HandModel.cal_distance, hand_model.py:875-987): dist plus what the backward needs,
 * arg_link (Bt,n_pts) int32 = the max link, dgrad (Bt,n_pts,3) = d(dist)/d(link-frame point).  The backward
 * ACCUMULATES d(loss)/d(link_tf) for a cotangent grad_dist (Bt,n_pts); the points themselves get no gradient (the
 * reference passes constant surface samples, energy.py:57-59). */
int gqp_hand_distance_full(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each, int n_pts,
                           int per_candidate_points, float *dist, int32_t *arg_link, float *dgrad, float *scratch,
                           void *stream);
int gqp_hand_distance_backward(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each,
                               int n_pts, int per_candidate_points, const float *grad_dist, const int32_t *arg_link,
                               const float *dgrad, float *grad_link_tf, void *stream);
uint64_t gqp_hand_distance_scratch_floats(int n_obj, int batch_each, int n_pts, int per_candidate_points);
/* The culled form (dist == NULL) with a BOUNDED survivor queue: queue_per_candidate samples per candidate ON AVERAGE over
 * the batch (0 = n_pts, the unbounded default of gqp_hand_distance: 20 KB of scratch per candidate at 2500 samples, where
 * 20-45 survivors per candidate are observed along a fit).  If the queue fills up, word 2 of the counters
 * (gqp_hand_distance_counter_offset) is set and e_pen / the gradient of THAT call are invalid: callers check it at
 * their next synchronisation point (FitEngine.assert_ok) — nothing is dropped silently. */
uint64_t gqp_hand_distance_scratch_floats_bounded(int n_obj, int batch_each, int n_pts, int per_candidate_points,
                                                  int queue_per_candidate);
int gqp_hand_distance_bounded(const GqpHand *h, const float *link_tf, const float *x, int n_obj, int batch_each,
                              int n_pts, int per_candidate_points, float *e_pen, float *scratch, float w_pen,
                              float *grad_link_tf, int queue_per_candidate, void *stream);

/* diagnostics: index (in 32-bit words) inside `scratch` of the culled path's counters after gqp_hand_distance(dist = NULL):
 * word 0 = samples that survived culling and were queued for the exact search, word 1 = (candidate, 32-sample chunk) work
 * items, word 2 = 1 if a bounded queue overflowed (gqp_hand_distance_bounded) */
uint64_t gqp_hand_distance_counter_offset(int n_obj, int batch_each, int n_pts, int per_candidate_points,
                                          const void *scratch);

/* diagnostics: info_host[6] = {shared-memory blob bytes, BVH nodes, sign-grid cells, cells proven positive, faces, grids} */
int gqp_hand_pen_info(const GqpHand *h, int32_t *info_host);

/* diagnostics for the culling-proof tests: the outside grid of mesh link `link` exactly as the kernels use it.
 * geom_host[10] = grid origin xyz, 1/cell, link box lo xyz, hi xyz (link frame); dim_host[3] = cells per axis, all 0 when
 * the link has no grid (mesh not proven "nice": never culled); bits_host (may be NULL) receives ceil(cells/32) words,
 * cell (x,y,z) -> bit (z*dim[1]+y)*dim[0]+x.  The kernels look a point q up as cell = (int)((q - origin) * (1/cell)).
 * A SET bit, or q outside [lo,hi], is the claim "the reference's sign is +1 for every point there". */
int gqp_hand_link_grid(const GqpHand *h, int link, float *geom_host, int32_t *dim_host, uint32_t *bits_host,
                       int max_words);

/* The link's "tie planes" (4 floats each: unit normal, offset) -> returns their number (>= 0) or a negative error.
 * slab_host[3] (may be NULL) = {NEG, POS, number of anchors}; planes_host then receives the planes followed by the anchor
 * vertices (xyz0), max_planes counting both.  A point q of the link frame with  -NEG*D <= n.q - offset <= POS*D  for one
 * of the planes is NEVER culled, first with D = |q - box centre| + box half-diagonal, then (points only) with
 * D = 1.001 * distance to the nearest anchor: behind the extended plane of a face at a right-angle crease the reference's
 * fp32 first-minimum rule can return sign -1 for outside points (csrc/gqp_bvh.cpp: gqp_mesh_tie_planes). */
int gqp_hand_link_planes(const GqpHand *h, int link, float *planes_host, int max_planes, float *slab_host);

/* ---- H5 + E_joints: self_penetration (hand_model.py:989-1040) and joint limits (energy.py:47-54) ------------------
 * e_spen (Bt), e_joints (Bt); ACCUMULATES w_spen*d(e_spen) into grad_link_tf and OVERWRITES grad_pose_joints
 * (Bt, 9+n_dofs) with w_joints*d(e_joints)/d(pose) (both may be NULL). */
int gqp_self_pen_joints(const GqpHand *h, const float *link_tf, const float *pose, int Bt, float *e_spen,
                        float *e_joints, float w_spen, float w_joints, float *grad_link_tf, float *grad_pose_joints,
                        void *stream);

/* ---- (f3) the optional terms E_prior / E_wall (core/energy.py:68-76; --w_prior / --w_wall, both 0 by default) --------------
 * e_prior (Bt) = 1 - (R_g grasp_axis) . (0,0,-1), grasp_axis_host[3] = HandModel.grasp_axis (hand_model.py:470);
 * e_wall (Bt) = sum_i |min(z_i, 0)| over the n_surf hand surface samples `surf` (n_surf,3, device, link frames) of links
 * `surf_link` (n_surf, device) — HandModel.get_surface_points (hand_model.py:1042-1071).  ACCUMULATES w_wall*d(e_wall) into
 * grad_link_tf (atomics) and ADDS w_prior*d(e_prior)/d(pose[:,3:9]) to grad_pose (Bt, 9+n_dofs); both may be NULL.
 * Call it after gqp_self_pen_joints (which overwrites grad_pose) and before gqp_fk_contacts_backward.  surf_link entries must
 * be valid link ids.  E_manipulativity (energy.py:78-87) is not offered: scripts/fit.py's weight_dict has no entry for it. */
int gqp_prior_wall(const GqpHand *h, const float *link_tf, const float *pose, int Bt, const float *grasp_axis_host,
                   const float *surf, const int32_t *surf_link, int n_surf, float w_prior, float w_wall, float *e_prior,
                   float *e_wall, float *grad_link_tf, float *grad_pose, void *stream);

/* ---- Q1-Q4 + E_dis: force-closure QP (span.py:263-415, qp_solver.py:60-134, registry.py:31-89) and contact-distance
 * energy (energy.py:25-28), forward + analytic backward in one launch, a warp per candidate. -----------------------
 * cpts/cnrm_hand (Bt,n_c,3) from gqp_contacts_forward; obj_dist (Bt,n_c), obj_nrm (Bt,n_c,3) from
 * gqp_object_distance_forward; obj_dgrad (Bt,n_c,3) = d(obj_dist)/d(cpts) (gqp_object_distance_backward with ones);
 * cog (n_obj,3).  Outputs e_fc, e_dis (Bt), lam (Bt,n_c) per-contact force sums (with_solution=True), and, if not
 * NULL, grad_cpts / grad_cnrm (Bt,n_c,3) = d(w_fc*e_fc + w_dis*e_dis)/d(.), OVERWRITTEN.
 * qp_stop_iter: NULL = run max_iter predictor-corrector iterations; else device int: leave the loop at that iteration
 * (emulates qpth's batch-wide exit test); qp_trace (Bt,max_iter) best-residual trace or NULL. */
typedef struct GqpFcParams {
    float friction, torque_weight, lower, upper, ridge, svd_gain, values_gain, values_offset, eps;
    int32_t max_iter;
    float w_fc, w_dis;
} GqpFcParams;
int gqp_fc_dis_energy(const GqpFcParams *p_host, const float *cpts, const float *cnrm_hand, const float *obj_dist,
                      const float *obj_nrm, const float *obj_dgrad, const float *cog, int n_obj, int batch_each,
                      int n_c, float *e_fc, float *e_dis, float *lam, float *grad_cpts, float *grad_cnrm,
                      const int32_t *qp_stop_iter, float *qp_trace, void *stream);
/* batch-wide exit iteration of qpth's loop from the trace (one block): writes *stop_iter. */
int gqp_fc_stop_iter(const float *qp_trace, int Bt, int max_iter, float eps, int32_t *stop_iter, void *stream);

/* The same computation split so that qpth's batch-wide exit test costs ONE interior-point pass instead of two (what
 * the engine launches every iteration):
 *   gqp_fc_solve  runs max_iter iterations per candidate; into `workspace` (gqp_fc_workspace_bytes bytes, 8-byte
 *     aligned, caller-owned) it stores every iterate that improved the candidate's best residual, which iterate was the
 *     best as of each iteration, and per iteration the batch maximum of the best residuals and an any-improved flag;
 *   gqp_fc_finish derives the exit iteration from those flags (batch_stop != 0; else every candidate keeps its best
 *     of all max_iter iterates), then computes value, svd term, E_fc, E_dis, lam and the analytic backward from each
 *     candidate's stored best iterate as of the exit iteration.  *stop_iter_out (device, may be NULL) gets the exit
 *     iteration.  Outputs as gqp_fc_dis_energy. */
uint64_t gqp_fc_workspace_bytes(int Bt, int n_c, int max_iter);
int gqp_fc_solve(const GqpFcParams *p_host, const float *cpts, const float *obj_nrm, const float *cog, int n_obj,
                 int batch_each, int n_c, void *workspace, void *stream);
int gqp_fc_finish(const GqpFcParams *p_host, const float *cpts, const float *cnrm_hand, const float *obj_dist,
                  const float *obj_nrm, const float *obj_dgrad, const float *cog, int n_obj, int batch_each, int n_c,
                  const void *workspace, int batch_stop, float *e_fc, float *e_dis, float *lam, float *grad_cpts,
                  float *grad_cnrm, int32_t *stop_iter_out, void *stream);

/* ---- (f3) the reference's other force-closure energies (scripts/fit.py:344-347) + E_dis, forward and analytic backward in
 * one launch, a warp per candidate; replaces gqp_fc_solve + gqp_fc_finish for energy_type dexgrasp / tdg. ---------------
 *   GQP_SPAN_DEXGRASP  metrics/ops/dexgrasp.py:5-69: e_fc = |sum n_i|^2 + tw^2 |sum n_i x (p_i - cog)|^2; torque_weight is
 *                      0.0 on the reference's path (DexgraspSpanMetric.forward's default; energy.py:34-41 passes none).
 *   GQP_SPAN_TDG       metrics/ops/tdg.py:57-239: e_fc = scale * mean over the n_dirs unit directions `dirs` (n_dirs,3)
 *                      (device; TDGEnergy draws 1000 with np.random.randn, :44-48, 148-150) of 1 - cos(W(u), [u;0]);
 *                      friction = miu_coef[0] = 0.2, obb_length = 0.2, scale = 100, enable_density = 1 (:219-227, :236).
 * Inputs and outputs as gqp_fc_finish (lam is not produced: the reference returns ones / None there).  n_c <= 32. */
#define GQP_SPAN_DEXGRASP 1
#define GQP_SPAN_TDG 2
typedef struct GqpSpanParams {
    int32_t kind;
    float torque_weight;                 /* dexgrasp */
    float friction, obb_length, scale;   /* tdg */
    int32_t enable_density;              /* tdg */
    float w_fc, w_dis;
} GqpSpanParams;
int gqp_span_energy(const GqpSpanParams *p_host, const float *cpts, const float *cnrm_hand, const float *obj_dist,
                    const float *obj_nrm, const float *obj_dgrad, const float *cog, const float *dirs, int n_dirs,
                    int n_obj, int batch_each, int n_c, float *e_fc, float *e_dis, float *grad_cpts, float *grad_cnrm,
                    void *stream);

/* ---- M1/M2/M4: MALA* (optimizer.py:199-340) and the caller's bookkeeping (fit.py:403-406, 434-458) -----------------
 * State rows are candidates.  u_switch (Bt,n_c), idx_rand (Bt,n_c) int32, u_accept (Bt): injected noise, or all NULL
 * to draw them from Philox4x32-10 keyed (seed, iteration, row).
 * gqp_mala_stats: grad2_mean (D) = mean over ALL rows of grad^2 (optimizer.py:231) and the per-object z-score of
 *   `energy` (fit.py:403-406, unbiased std) -> z (Bt).   Deterministic two-pass reduction.
 * gqp_mala_propose: ema = mu*grad2_mean + (1-mu)*ema (NaN->0); pose_new = pose - s_i*grad/(sqrt(ema)+1e-6);
 *   with clip_grad NaN gradients count as 0 and the rest is clamped to +-100 (:211-213); a row of pose_new with a NaN
 *   entry is zeroed as a whole (:242-244);
 *   idx_new = switch ? idx_rand : idx;  step += 1; step_size_out (Bt) = s_i (may be NULL).
 * gqp_mala_accept: new_energy = sum_k w_k*losses_new[k]; T = T0*decay^(step/period)*(1+Phi(z));
 *   accept = u < exp((E-E')/T) or reset; accepted rows copy pose_new/idx_new/grad_new/energy/losses into the state.
 *   after_reset != 0 (pass it in the iteration that FOLLOWS a reset iteration): rejected rows keep grad + grad_new instead of
 *   grad.  This restates what the reference does there: its proposals are not detached (optimizer.py:240, hand_model.py:855),
 *   the pose a reset iteration builds is an autograd leaf (hand_model.py:845-850) whose .grad accumulates in place, and
 *   old_grad_hand_pose (:268) aliases it — so :338 hands rejected rows old + new gradient.  Pinned by the reference run in
 *   tests/golden/fit_alt_ref.npz. */
typedef struct GqpMalaParams {
    float switch_possibility, starting_temperature, temperature_decay, step_size, mu;
    int32_t annealing_period, stepsize_period;
    int32_t clip_grad;
} GqpMalaParams;
int gqp_mala_stats(const float *grad, const float *energy, int n_obj, int batch_each, int D, int clip_grad,
                   float *grad2_mean, float *z, float *scratch /* ceil(Bt/64)*D floats */, void *stream);
/* gqp_mala_stats without caller-owned scratch and without ANY allocation (no cudaMalloc / cudaFree / allocating library
 * call): the scratch is a static 256 KB device array inside the library (gqp_mala_stats_noalloc_max_floats() = 65536
 * floats; needed: ceil(Bt/64)*D, e.g. 102 400 Allegro candidates), held by every GPU that loads the library until exit.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate (CUDA
 * graph capture, serving loops).  Sizes beyond the maximum return GQP_ERR_LIMIT.  All calls share the array (it keeps the
 * last call's contents; nothing in it needs to start at zero): calls that OVERLAP on the GPU (different streams) race.
 * Same results as gqp_mala_stats, bit for bit. */
int gqp_mala_stats_noalloc(const float *grad, const float *energy, int n_obj, int batch_each, int D, int clip_grad,
                           float *grad2_mean, float *z, void *stream);
uint64_t gqp_mala_stats_noalloc_max_floats(void);
int gqp_mala_propose(const GqpMalaParams *p_host, const float *pose, const float *grad, const float *grad2_mean,
                     float *ema, int32_t *step, const int32_t *idx, int Bt, int D, int n_c, int n_cand,
                     const float *u_switch, const int32_t *idx_rand, uint64_t seed, uint64_t iteration, float *pose_new,
                     int32_t *idx_new, float *step_size_out, void *stream);
int gqp_mala_accept(const GqpMalaParams *p_host, int Bt, int D, int n_c, int n_terms, const float *weights /* device (n_terms) */,
                    const float *losses_new /* (n_terms,Bt) */, const float *z, const uint8_t *reset_mask,
                    const float *u_accept, uint64_t seed, uint64_t iteration, const int32_t *step,
                    const float *pose_new, const int32_t *idx_new, const float *grad_new, float *pose, int32_t *idx,
                    float *grad, float *energy, float *losses /* (n_terms,Bt) */, float *new_energy_out,
                    uint8_t *accept_out, float *temperature_out, int after_reset, void *stream);

/* ---- (f3) AnnealingDexGraspNet (core/optimizer.py:11-149; scripts/fit.py --optimizer dexgraspnet) --------------------------
 * One RMS accumulator ema (D) and ONE step counter for the whole batch (passed by value: the reference keeps it as a host
 * int), no clipping, no NaN handling, temperature without the z-score factor, resets do not touch the optimiser.
 * gqp_anneal_propose: step = the counter BEFORE this try_step (:81); ema_out = mu*grad2_mean + (1-mu)*ema_in (different
 *   buffers: the caller swaps them); pose_new = pose - s*grad/(sqrt(ema_out)+1e-6); idx_new as gqp_mala_propose;
 *   *step_size_out (one device float, may be NULL) = s.  grad2_mean from gqp_mala_stats (clip_grad = 0).
 * gqp_anneal_accept: step = the counter AFTER the try_step (:119-121); accept = u < exp((E-E')/T); commit as
 *   gqp_mala_accept (after_reset included).  temperature_out (Bt) or NULL. */
int gqp_anneal_propose(const GqpMalaParams *p_host, const float *pose, const float *grad, const float *grad2_mean,
                       const float *ema_in, float *ema_out, int step, const int32_t *idx, int Bt, int D, int n_c,
                       int n_cand, const float *u_switch, const int32_t *idx_rand, uint64_t seed, uint64_t iteration,
                       float *pose_new, int32_t *idx_new, float *step_size_out, void *stream);
int gqp_anneal_accept(const GqpMalaParams *p_host, int Bt, int D, int n_c, int n_terms, const float *weights,
                      const float *losses_new, const float *u_accept, uint64_t seed, uint64_t iteration, int step,
                      const float *pose_new, const int32_t *idx_new, const float *grad_new, float *pose, int32_t *idx,
                      float *grad, float *energy, float *losses, float *new_energy_out, uint8_t *accept_out,
                      float *temperature_out, int after_reset, void *stream);

/* ---- (f4) entropy statistics of the batch (HandModel.joint_entropy / pose_entropy, hand_model.py:698-760; --log_entropy) ----
 * 32-bin histograms over ALL Bt rows (torch.histc semantics): every joint between its limits (joints_lower / joints_upper:
 * DEVICE arrays of n_dofs), translation components in [-0.1, 0.1], rotation as (r, theta, phi) of its rotation vector.
 * out3 (device): [joints_entropy, translation_entropy, rotation_entropy]; scratch: n_dofs + 6 floats (per-histogram
 * entropies).  roma.rotmat_to_rotvec is restated (w >= 0 quaternion): PARITY UNPINNED for the rotation part. */
int gqp_entropy_stats(const float *pose, int Bt, int n_dofs, const float *joints_lower, const float *joints_upper,
                      float *scratch, float *out3, void *stream);

/* ---- M3: outlier reset (fit.py:408-422: reset_mask = z > threshold; initialize_convex_hull(env_mask); reset_envs) ----
 * Called between gqp_mala_propose and the energy evaluation of a reset iteration.  pose_init (Bt,D) / idx_init (Bt,n_c)
 * hold a fresh initialisation of EVERY row (initialize_convex_hull recomputes all rows, core/initializations.py:15-193:
 * gqp_init_convex_hull below, or injected).  mask = mask_in (uint8, injected) or, if NULL, z > z_threshold.  Masked rows:
 * pose_new <- pose_init, idx_new <- idx_init (HandModel.set_parameters(env_mask), hand_model.py:845-850, 816-822),
 * step <- 0, ema <- 0 (MalaStar.reset_envs, optimizer.py:275-287).  mask_out (Bt) uint8 is what gqp_mala_accept takes as
 * reset_mask (forced accept, optimizer.py:315-316); *count_out (device int, may be NULL) = number of masked rows.
 * idx_eval (Bt,n_c), may be NULL: receives idx_init for EVERY row — the reference gathers the contact points of a reset
 * iteration with the freshly drawn indices for all rows (hand_model.py:830-831 uses the argument, not the merged
 * member), so that iteration's energies are evaluated at idx_eval while idx_new is what gets committed. */
int gqp_mala_reset(int Bt, int D, int n_c, const float *z, float z_threshold, const uint8_t *mask_in,
                   const float *pose_init, const int32_t *idx_init, float *pose_new, int32_t *idx_new,
                   int32_t *idx_eval, int32_t *step, float *ema, uint8_t *mask_out, int32_t *count_out, void *stream);

/* ---- (f1) initialize_convex_hull on the device (core/initializations.py:15-193; the reference runs it on the host with
 * trimesh + pytorch3d for ALL objects at every reset, scripts/fit.py:421) --------------------------------------------------
 * gqp_init_create: hull_tris_host (H_total,3,3) outward-oriented convex-hull triangles of the objects, object o owns faces
 *   [hull_offset_host[o], hull_offset_host[o+1]) (the hull itself is setup-time host work, as trimesh's .convex_hull).
 * gqp_init_convex_hull: for every object pool_factor*batch_each area-uniform hull samples pushed `inflate` along their
 *   face normal, farthest point sampling of batch_each of them (start index 0), then per candidate the pose
 *   [p - distance*n, first two columns of look_at(p, p+n) @ euler_rxyz(tilt, pitch, rotate), truncated-normal joints]
 *   and n_c contact indices.  Noise: u_pool (n_obj, pool_factor*batch_each, 3), u_pose (Bt, 4+n_dofs) uniforms in [0,1)
 *   (distance, rotate, pitch, tilt, joints) and idx_rand (Bt,n_c) injected, or NULL = Philox4x32-10 keyed (seed, counter).
 *   pose_init (Bt, 9+n_dofs), idx_init (Bt,n_c) are what gqp_mala_reset consumes; picked_out (Bt) int32, may be NULL,
 *   receives the pool index picked for every candidate.  scratch: gqp_init_scratch_bytes bytes, 16-byte aligned.
 *   Not restated: trimesh.sample.sample_surface_even's removal of samples closer than a radius before the FPS (the FPS
 *   does the spreading). */
typedef struct GqpInit GqpInit;
typedef struct GqpInitParams {
    float distance_lower, distance_upper, rotate_lower, rotate_upper, pitch_lower, pitch_upper, tilt_lower, tilt_upper;
    float jitter_strength, inflate;
    int32_t pool_factor;
    float forward_axis[3], up_axis[3]; /* HandModel.forward_axis / up_axis (hand_model.py:459-468) */
    float default_state[32];           /* HandModel.default_state (n_dofs entries) */
} GqpInitParams;
int gqp_init_create(const float *hull_tris_host, const int32_t *hull_offset_host, int n_obj, GqpInit **out);
void gqp_init_destroy(GqpInit *init);
uint64_t gqp_init_scratch_bytes(int n_obj, int batch_each, int pool_factor);
int gqp_init_convex_hull(const GqpInit *init, const GqpHand *h, const GqpInitParams *p_host, int batch_each, int n_c,
                         const float *u_pool, const float *u_pose, const int32_t *idx_rand, uint64_t seed,
                         uint64_t counter, void *scratch, float *pose_init, int32_t *idx_init, int32_t *picked_out,
                         void *stream);

/* cudaMemsetAsync(ptr, 0, bytes) on the stream: the per-step reset of grad_link_tf. */
int gqp_memset_zero(void *ptr, uint64_t bytes, void *stream);

/* FP32 FMA micro-benchmark (roofline denominator for the SIMT stages): runs `iters` dependent-chain FMAs on every
 * lane of a full-chip grid; returns the FMA count in *n_fma_host.  Timed by the caller with CUDA events. */
int gqp_fma_peak_probe(int iters, float *sink, uint64_t *n_fma_host, void *stream);
/* the same for FP64 (the QP stage computes in double): roofline denominator of fc_dis_kernel */
int gqp_dfma_peak_probe(int iters, double *sink, uint64_t *n_fma_host, void *stream);

#ifdef __cplusplus
}
#endif
#endif
