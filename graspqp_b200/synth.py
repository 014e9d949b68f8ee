"""Synthetic inputs for tests and benchmarks (the reference ships no dataset: scripts/fit.py:124-125).

Procedural closed, outward-oriented triangle meshes (sphere, box, cylinder, torus, random convex hulls), area-weighted
surface samples thinned by farthest-point sampling (the reference samples 100x then FPS, object_model.py:166-171),
and random-init grasp poses in the spirit of core/initializations.py:100-193 (approach along the inward normal at
distance U(0.05,0.1) from a surface point (fit.py:61-62), joints = truncated normal around default_state).
Host-side numpy only; nothing here is on the timed path.
"""
import numpy as np


def _finish(verts, faces, radius):
    verts = np.asarray(verts, np.float64)
    verts = verts - verts.mean(0)
    verts = verts * (radius / np.linalg.norm(verts, axis=1).max())
    return verts.astype(np.float32), np.asarray(faces, np.int64)


def uv_sphere(n_faces=2000, radius=0.08):
    nlat = max(3, int(round(np.sqrt(n_faces / 4.0))))
    nlon = max(3, int(round(n_faces / (2.0 * (nlat - 1)))))
    v = [[0, 0, 1.0]]
    for i in range(1, nlat):
        th = np.pi * i / nlat
        for j in range(nlon):
            ph = 2 * np.pi * j / nlon
            v.append([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
    v.append([0, 0, -1.0])
    f = []
    ring = lambda i, j: 1 + (i - 1) * nlon + (j % nlon)
    for j in range(nlon):
        f.append([0, ring(1, j), ring(1, j + 1)])
        f.append([len(v) - 1, ring(nlat - 1, j + 1), ring(nlat - 1, j)])
    for i in range(1, nlat - 1):
        for j in range(nlon):
            a, b, c, d = ring(i, j), ring(i, j + 1), ring(i + 1, j), ring(i + 1, j + 1)
            f.append([a, c, d])
            f.append([a, d, b])
    return _finish(v, f, radius)


def _grid_surface(fn, nu, nv, wrap_u, wrap_v):
    """Triangulate a parametric grid fn(u,v)->xyz, u in [0,1), v in [0,1)."""
    U = nu if wrap_u else nu + 1
    V = nv if wrap_v else nv + 1
    v = [fn(i / nu, j / nv) for i in range(U) for j in range(V)]
    idx = lambda i, j: (i % U) * V + (j % V)
    f = []
    for i in range(nu):
        for j in range(nv):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i, j + 1), idx(i + 1, j + 1)
            f.append([a, b, d])
            f.append([a, d, c])
    return v, f


def torus(n_faces=10000, radius=0.09, tube=0.35):
    nu = max(3, int(round(np.sqrt(n_faces / 2.0 / tube))))
    nv = max(3, int(round(n_faces / 2.0 / nu)))

    def fn(u, v):
        a, b = 2 * np.pi * u, 2 * np.pi * v
        return [(1 + tube * np.cos(b)) * np.cos(a), (1 + tube * np.cos(b)) * np.sin(a), tube * np.sin(b)]

    v, f = _grid_surface(fn, nu, nv, True, True)
    return _finish(v, f, radius)


def box(n_faces=10000, radius=0.09, extents=(1.0, 0.7, 0.5)):
    n = max(1, int(round(np.sqrt(n_faces / 12.0))))
    verts, faces, cache = [], [], {}

    def vid(p):
        k = tuple(np.round(p, 9))
        if k not in cache:
            cache[k] = len(verts)
            verts.append(list(p))
        return cache[k]

    e = np.asarray(extents) / 2
    for ax in range(3):
        for sgn in (-1, 1):
            u_ax, v_ax = (ax + 1) % 3, (ax + 2) % 3
            for i in range(n):
                for j in range(n):
                    q = []
                    for di, dj in ((0, 0), (1, 0), (1, 1), (0, 1)):
                        p = np.zeros(3)
                        p[ax] = sgn * e[ax]
                        p[u_ax] = (-1 + 2 * (i + di) / n) * e[u_ax]
                        p[v_ax] = (-1 + 2 * (j + dj) / n) * e[v_ax]
                        q.append(vid(p))
                    if sgn > 0:
                        faces += [[q[0], q[1], q[2]], [q[0], q[2], q[3]]]
                    else:
                        faces += [[q[0], q[2], q[1]], [q[0], q[3], q[2]]]
    return _finish(verts, faces, radius)


def cylinder(n_faces=10000, radius=0.09, aspect=1.2):
    nseg = max(3, int(round(np.sqrt(n_faces / 4.0))))
    nh = max(1, int(round((n_faces - 2 * nseg) / (2.0 * nseg))))
    v, f = [], []
    for k in range(nh + 1):
        z = aspect * (-1 + 2 * k / nh)
        for j in range(nseg):
            v.append([np.cos(2 * np.pi * j / nseg), np.sin(2 * np.pi * j / nseg), z])
    idx = lambda k, j: k * nseg + (j % nseg)
    for k in range(nh):
        for j in range(nseg):
            a, b, c, d = idx(k, j), idx(k, j + 1), idx(k + 1, j), idx(k + 1, j + 1)
            f += [[a, b, d], [a, d, c]]
    bot, top = len(v), len(v) + 1
    v += [[0, 0, -aspect], [0, 0, aspect]]
    for j in range(nseg):
        f.append([bot, idx(0, j + 1), idx(0, j)])
        f.append([top, idx(nh, j), idx(nh, j + 1)])
    return _finish(v, f, radius)


def convex_hull(n_faces=10000, radius=0.09, seed=0):
    from scipy.spatial import ConvexHull

    rng = np.random.default_rng(seed)
    n_pts = n_faces // 2 + 2  # a triangulated sphere-like hull has F = 2V - 4
    p = rng.normal(size=(n_pts, 3))
    p = p / np.linalg.norm(p, axis=1, keepdims=True) * rng.uniform(0.6, 1.0, size=3)
    hull = ConvexHull(p)
    f = hull.simplices.copy()
    c = p.mean(0)
    n = np.cross(p[f[:, 1]] - p[f[:, 0]], p[f[:, 2]] - p[f[:, 0]])
    flip = ((p[f[:, 0]] - c) * n).sum(1) < 0
    f[flip] = f[flip][:, [0, 2, 1]]
    used = np.unique(f)
    remap = -np.ones(len(p), np.int64)
    remap[used] = np.arange(len(used))
    return _finish(p[used], remap[f], radius)


def make_mesh(kind, n_faces, seed=0):
    radius = 0.06 + 0.06 * ((seed * 0.6180339887) % 1.0)  # 0.06 .. 0.12 m
    if kind == "sphere":
        return uv_sphere(n_faces, radius)
    if kind == "box":
        return box(n_faces, radius)
    if kind == "cylinder":
        return cylinder(n_faces, radius)
    if kind == "torus":
        return torus(n_faces, radius)
    if kind == "hull":
        return convex_hull(n_faces, radius, seed)
    raise ValueError(kind)


def face_vertices(verts, faces):
    """index_vertices_by_faces (thirdparty/TorchSDF/torchsdf/sdf.py:5-31) on the host."""
    return np.ascontiguousarray(verts[faces], np.float32)


def sample_surface(verts, faces, n, seed=1, oversample=4):
    rng = np.random.default_rng(seed)
    t = verts[faces].astype(np.float64)
    area = 0.5 * np.linalg.norm(np.cross(t[:, 1] - t[:, 0], t[:, 2] - t[:, 0]), axis=1)
    m = n * oversample
    fi = rng.choice(len(faces), size=m, p=area / area.sum())
    u, v = rng.random(m), rng.random(m)
    su = np.sqrt(u)
    w = np.stack([1 - su, su * (1 - v), su * v], 1)
    pts = (t[fi] * w[:, :, None]).sum(1)
    # farthest point sampling
    sel = np.empty(n, np.int64)
    sel[0] = 0
    d = np.linalg.norm(pts - pts[0], axis=1)
    for k in range(1, n):
        sel[k] = int(d.argmax())
        d = np.minimum(d, np.linalg.norm(pts - pts[sel[k]], axis=1))
    return pts[sel].astype(np.float32)


def random_grasp_init(pack, verts_list, batch_each, n_contact=12, seed=1, jitter=0.1):
    """Returns hand_pose (O*B, 9+dof) float32 and contact indices (O*B, n_contact) int64."""
    rng = np.random.default_rng(seed)
    lo, hi, mu = pack["joints_lower"], pack["joints_upper"], pack["default_state"]
    dof = len(lo)
    poses = []
    for verts in verts_list:
        B = batch_each
        p = verts[rng.integers(0, len(verts), B)].astype(np.float64)
        c = verts.mean(0)
        n = p - c
        n /= np.linalg.norm(n, axis=1, keepdims=True) + 1e-12
        t = p + n * rng.uniform(0.05, 0.1, (B, 1))
        # hand "forward" (+z for allegro) looks at the object, random roll about it
        z = -n
        a = rng.normal(size=(B, 3))
        x = a - (a * z).sum(1, keepdims=True) * z
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        y = np.cross(z, x)
        R = np.stack([x, y, z], -1)  # columns
        six = np.concatenate([R[:, :, 0], R[:, :, 1]], 1)
        sig = jitter * (hi - lo)
        th = np.clip(mu + sig * np.clip(rng.normal(size=(B, dof)), -2, 2), lo, hi)
        poses.append(np.concatenate([t, six, th], 1))
    pose = np.concatenate(poses, 0).astype(np.float32)
    idx = rng.integers(0, len(pack["cand"]), (len(pose), n_contact)).astype(np.int64)
    return pose, idx
