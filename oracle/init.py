"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Restatement of graspqp/src/graspqp/core/initializations.py:15-193 (`initialize_convex_hull`) with INJECTED noise, for
the device re-initialisation used by the outlier resets (scripts/fit.py:408-422).

What the reference does per object                                     (file:line)            here
  inflated convex hull of the mesh (trimesh -> qhull)                   :37-46                 scipy.spatial.ConvexHull
  100*B surface samples, `trimesh.sample.sample_surface_even`           :54                    area-weighted face + uniform
      (area-weighted + removal of samples closer than a radius)                                barycentric point; NO thinning
  points pushed 1 cm along their face normal                            :56                    same
  farthest point sampling of B of them, start index 0 (pytorch3d)       :60-62                 fps()
  n = unit(closest point on the hull - p): the inward face normal       :63-79                 -face normal (convex hull)
  rotation_global = look_at(p, p + n, up, forward)                      :84-120                look_at()
  distance, rotate, pitch, tilt ~ U(lower, upper)                       :122-136               injected uniforms
  rotation_local = transforms3d euler2mat(tilt, pitch, rotate, 'rxyz')  :138-145               Rx(tilt) Ry(pitch) Rz(rotate)
  translation = p - distance * n, rotation = global @ local             :147-150               same
  joints ~ torch.nn.init.trunc_normal_(mu, 0.1*range, lo-1e-6, hi+1e-6) :164-176               trunc_normal()
  hand_pose = [t, first two rows of R^T, joints], randint contacts      :187-193               same

PARITY UNPINNED for the pieces whose packages are absent from this image and from /root/reference: trimesh (hull, even
sampling, nearest.on_surface), pytorch3d (FPS), transforms3d (euler2mat): restated from their documented behaviour.
`trunc_normal` follows torch.nn.init.trunc_normal_ (checked against torch itself in tests/test_init.py); `look_at` is a
line-by-line restatement of the reference's own nested function.  All arithmetic is float32 in the kernel's operation
order so that the device kernel can be compared tightly (FPS indices bit-exact).
"""
import math

import numpy as np

F32 = np.float32


def convex_hull(vertices):
    """-> (H,3,3) outward-oriented hull triangles (float32).  trimesh.Trimesh.convex_hull calls qhull as well."""
    from scipy.spatial import ConvexHull
    v = np.asarray(vertices, np.float64)
    hull = ConvexHull(v)
    tris = v[hull.simplices]  # (H,3,3)
    n = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    flip = (n * hull.equations[:, :3]).sum(1) < 0  # qhull's equations carry the outward normal
    tris[flip] = tris[flip][:, [0, 2, 1]]
    return np.ascontiguousarray(tris, F32)


def hull_tables(tris):
    """unit normals (H,3) and the cumulative area distribution (H) of the hull faces, float32 from float64 sums"""
    t = np.asarray(tris, np.float64)
    n = np.cross(t[:, 1] - t[:, 0], t[:, 2] - t[:, 0])
    a = np.linalg.norm(n, axis=1)
    cdf = np.cumsum(0.5 * a)
    return (n / a[:, None]).astype(F32), (cdf / cdf[-1]).astype(F32)


def sample_pool(tris, normals, cdf, u, inflate=0.01):
    """u (M,3) uniforms -> pool points (M,3) float32 pushed `inflate` along the face normal, and their faces (M)"""
    u = np.asarray(u, F32)
    f = np.minimum(np.searchsorted(cdf, u[:, 0], side="right"), len(cdf) - 1)  # first face with cdf > u
    s = np.sqrt(u[:, 1])
    w0, w1, w2 = F32(1.0) - s, s * (F32(1.0) - u[:, 2]), s * u[:, 2]
    a, b, c = tris[f, 0], tris[f, 1], tris[f, 2]
    p = (w0[:, None] * a + w1[:, None] * b) + w2[:, None] * c
    return (p + F32(inflate) * normals[f]).astype(F32), f.astype(np.int32)


def fps(points, K):
    """pytorch3d.ops.sample_farthest_points(points[None], K=K, random_start_point=False): indices (K), first maximum"""
    p = np.asarray(points, F32)
    sel = np.zeros(K, np.int64)
    mind = np.full(len(p), np.inf, F32)
    cur = 0
    for k in range(K):
        sel[k] = cur
        d = p - p[cur]
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        mind = np.minimum(mind, d2)
        cur = int(np.argmax(mind))
    return sel


def _unit(v):
    return v / np.sqrt((v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2])[:, None]


def look_at(camera, target, forward_vector, up_vector):
    """initializations.py:84-118, float32"""
    base_up = np.asarray(up_vector, F32)
    fwd_vec = np.asarray(forward_vector, F32)
    forward = _unit((camera - target).astype(F32))
    prod = (forward * base_up).sum(1, dtype=F32)
    up = np.where((np.abs(prod) < F32(0.95))[:, None], base_up[None], np.array([0, 1, 0], F32)[None]).astype(F32)
    right = _unit(np.cross(up, forward).astype(F32))
    upv = np.cross(forward, right).astype(F32)
    orient = np.stack([forward, upv, right], -1)  # columns
    basis = np.stack([fwd_vec, -np.cross(fwd_vec, base_up), base_up], -1).astype(F32)
    return (orient @ basis).astype(F32)


def euler_rxyz(ai, aj, ak):
    """transforms3d.euler.euler2mat(ai, aj, ak, axes='rxyz'): rotating-frame x, y', z'' = Rx(ai) @ Ry(aj) @ Rz(ak)"""
    ci, si, cj, sj, ck, sk = np.cos(ai), np.sin(ai), np.cos(aj), np.sin(aj), np.cos(ak), np.sin(ak)
    R = np.empty((len(ai), 3, 3), F32)
    R[:, 0, 0], R[:, 0, 1], R[:, 0, 2] = cj * ck, -cj * sk, sj
    R[:, 1, 0], R[:, 1, 1], R[:, 1, 2] = si * sj * ck + ci * sk, -si * sj * sk + ci * ck, -si * cj
    R[:, 2, 0], R[:, 2, 1], R[:, 2, 2] = -ci * sj * ck + si * sk, ci * sj * sk + si * ck, ci * cj
    return R


def trunc_normal(u, mean, std, a, b):
    """torch.nn.init.trunc_normal_ with the uniform draw u in [0,1) injected (torch/nn/init.py: _no_grad_trunc_normal_)"""
    from scipy.special import erfinv
    ncdf = lambda x: (1.0 + math.erf(x / math.sqrt(2.0))) / 2.0
    lo, hi = 2 * ncdf((a - mean) / std) - 1, 2 * ncdf((b - mean) / std) - 1
    x = np.asarray(u, F32) * F32(hi - lo) + F32(lo)
    x = erfinv(x.astype(np.float64)).astype(F32) * F32(std * math.sqrt(2.0)) + F32(mean)
    return np.clip(x, F32(a), F32(b)).astype(F32)


DEFAULTS = dict(distance_lower=0.05, distance_upper=0.1, rotate_lower=-math.pi, rotate_upper=math.pi,
                pitch_lower=-15 * math.pi / 180, pitch_upper=15 * math.pi / 180, tilt_lower=-45 * math.pi / 180,
                tilt_upper=45 * math.pi / 180, jitter_strength=0.1, inflate=0.01, pool_factor=100)  # fit.py:59-71
AXIS = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1), "-x": (-1, 0, 0), "-y": (0, -1, 0), "-z": (0, 0, -1)}


def initialize_convex_hull(pack, hull_tris_list, batch_each, n_contact, u_pool, u_pose, idx_rand, **kw):
    """u_pool (O, pool_factor*B, 3), u_pose (O*B, 4+dof) uniforms in [0,1): distance, rotate, pitch, tilt, joints;
    idx_rand (O*B, n_contact).  -> hand_pose (O*B, 9+dof) float32, contact indices, and the FPS indices per object."""
    P = dict(DEFAULTS)
    P.update(kw)
    lo, hi = np.asarray(pack["joints_lower"], F32), np.asarray(pack["joints_upper"], F32)
    mu = np.clip(np.asarray(pack["default_state"], F32), lo, hi)
    dof, B = len(lo), batch_each
    fwd, up = AXIS[str(pack["axes"][0])], AXIS[str(pack["axes"][1])]
    poses, picks = [], []
    for o, tris in enumerate(hull_tris_list):
        normals, cdf = hull_tables(tris)
        pool, face = sample_pool(np.asarray(tris, F32), normals, cdf, u_pool[o], P["inflate"])
        sel = fps(pool, B)
        picks.append(sel)
        p, n = pool[sel], -normals[face[sel]]
        up_ = np.asarray(u_pose[o * B:(o + 1) * B], F32)
        lerp = lambda k, a: F32(P[a + "_lower"]) + F32(P[a + "_upper"] - P[a + "_lower"]) * up_[:, k]
        distance, rotate, pitch, tilt = lerp(0, "distance"), lerp(1, "rotate"), lerp(2, "pitch"), lerp(3, "tilt")
        Rg = look_at(p, (p + n).astype(F32), fwd, up)
        R = (Rg @ euler_rxyz(tilt, pitch, rotate)).astype(F32)
        t = (p - distance[:, None] * n).astype(F32)
        th = np.stack([trunc_normal(up_[:, 4 + j], float(mu[j]), float(P["jitter_strength"] * (hi[j] - lo[j])),
                                    float(lo[j]) - 1e-6, float(hi[j]) + 1e-6) for j in range(dof)], 1)
        six = np.concatenate([R[:, :, 0], R[:, :, 1]], 1)  # rotation.transpose(1,2)[:, :2].reshape(-1,6)
        poses.append(np.concatenate([t, six, th], 1).astype(F32))
    return np.concatenate(poses, 0), np.asarray(idx_rand).astype(np.int64), picks
