"""initialize_convex_hull — graspqp/src/graspqp/core/initializations.py:15-193 on the device (csrc/gqp_init.cu).

The reference computes, on the host and for every object at every reset (scripts/fit.py:421): the object's inflated convex
hull (trimesh), 100*B even surface samples, farthest point sampling of B of them (pytorch3d), a look-at rotation with Euler
jitter, a pose 5-10 cm off the hull and truncated-normal joint angles.  Here the hull is setup-time host work (scipy's
qhull, which is what trimesh calls) and everything per reset is three kernel launches.
"""
import ctypes
import math

import numpy as np

from . import _lib

AXIS = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1), "-x": (-1, 0, 0), "-y": (0, -1, 0), "-z": (0, 0, -1)}
# scripts/fit.py:59-71
DEFAULTS = dict(distance_lower=0.05, distance_upper=0.1, rotate_lower=-math.pi, rotate_upper=math.pi,
                pitch_lower=-15 * math.pi / 180, pitch_upper=15 * math.pi / 180, tilt_lower=-45 * math.pi / 180,
                tilt_upper=45 * math.pi / 180, jitter_strength=0.1, inflate=0.01, pool_factor=100)


def convex_hull(vertices):
    """(V,3) -> (H,3,3) float32 outward-oriented hull triangles (object_mesh.convex_hull, initializations.py:39)."""
    from scipy.spatial import ConvexHull
    v = np.asarray(vertices, np.float64)
    hull = ConvexHull(v)
    tris = v[hull.simplices]
    n = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    flip = (n * hull.equations[:, :3]).sum(1) < 0
    tris[flip] = tris[flip][:, [0, 2, 1]]
    return np.ascontiguousarray(tris, np.float32)


class DeviceInit:
    """Holds the hull tables of a group of objects on the device and launches gqp_init_convex_hull."""

    def __init__(self, cdll, backend, pack, hull_tris_list, batch_each, n_contact, **params):
        self.lib, self.be = cdll, backend
        self.n_obj, self.batch_each, self.n_c = len(hull_tris_list), int(batch_each), int(n_contact)
        P = dict(DEFAULTS)
        P.update(params)
        self.params = _lib.GqpInitParams(**{k: P[k] for k in DEFAULTS})
        self.params.forward_axis = (ctypes.c_float * 3)(*AXIS[str(pack["axes"][0])])
        self.params.up_axis = (ctypes.c_float * 3)(*AXIS[str(pack["axes"][1])])
        ds = np.zeros(32, np.float32)
        ds[:len(pack["default_state"])] = pack["default_state"]
        self.params.default_state = (ctypes.c_float * 32)(*ds.tolist())
        tris = np.ascontiguousarray(np.concatenate([np.asarray(t, np.float32).reshape(-1, 9) for t in hull_tris_list]))
        off = np.zeros(self.n_obj + 1, np.int32)
        off[1:] = np.cumsum([len(t) for t in hull_tris_list])
        self.handle = ctypes.c_void_p()
        _lib.check(cdll.gqp_init_create(tris.ctypes.data, off.ctypes.data, self.n_obj, ctypes.byref(self.handle)), cdll)
        nbytes = int(cdll.gqp_init_scratch_bytes(self.n_obj, self.batch_each, self.params.pool_factor))
        self.scratch = backend.empty(((nbytes + 15) // 16 * 2,), np.float64)  # 16-byte aligned by construction
        self.counter = 0

    def run(self, hand, pose_init, idx_init, u_pool=None, u_pose=None, idx_rand=None, seed=1, picked=None):
        """fills pose_init (Bt,D) / idx_init (Bt,n_c) on the backend's stream; noise injected or Philox (seed, call count)"""
        p = self.be.ptr
        self.counter += 1
        _lib.check(self.lib.gqp_init_convex_hull(self.handle, hand, ctypes.byref(self.params), self.batch_each, self.n_c,
                                                 p(u_pool), p(u_pose), p(idx_rand), int(seed), self.counter,
                                                 p(self.scratch), p(pose_init), p(idx_init), p(picked), self.be.stream),
                   self.lib)

    def close(self):
        if self.handle:
            self.lib.gqp_init_destroy(self.handle)
            self.handle = None
