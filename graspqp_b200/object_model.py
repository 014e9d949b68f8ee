"""ObjectModel — graspqp/src/graspqp/core/object_model.py for the hot path: object SDF at the contact points, surface
samples, centre of gravity.  Meshes come in as face-vertex arrays (the reference loads them with trimesh from
<data_root>/<code>/coacd/decomposed.obj, object_model.py:99-117; file IO is out of this path's scope)."""
import ctypes

import numpy as np
import torch

from . import _lib, ops

from . import SDF_BACKEND  # read from the environment like the reference (object_model.py:15-24); TORCHSDF contract


class ObjectModel:
    def __init__(self, data_root_path=None, batch_size_each=1, num_samples=2500, device="cuda", scale=1.0):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("graspqp_b200.ObjectModel runs on CUDA devices only (no CPU fallback)")
        self.batch_size_each = int(batch_size_each)
        self.num_samples = int(num_samples)
        self.object_scale = float(scale)  # object_model.py:60 — always 1.0 in the reference
        self.meshes, self.objset = [], None
        self.object_code_list = []

    def initialize(self, object_code_list, face_verts_list=None, surface_points=None, sdf_library=SDF_BACKEND):
        """object_model.py:70-184.  face_verts_list: list of (F,3,3) arrays; surface_points: (O, num_samples, 3)."""
        if str(sdf_library).upper() not in ("TORCHSDF", "B200"):
            raise RuntimeError(f"sdf_library={sdf_library!r}: only the TORCHSDF contract is implemented")
        if face_verts_list is None or surface_points is None:
            raise ValueError("pass face_verts_list and surface_points (mesh loading / sampling is setup-time host work)")
        lib = _lib.lib()
        self.object_code_list = list(object_code_list)
        with torch.cuda.device(self.device):
            for fv in face_verts_list:
                fv = np.ascontiguousarray(fv, np.float32).reshape(-1, 9)
                m = ctypes.c_void_p()
                _lib.check(lib.gqp_mesh_create(fv.ctypes.data, len(fv), ctypes.byref(m)))
                self.meshes.append(m)
            arr = (ctypes.c_void_p * len(self.meshes))(*[m.value for m in self.meshes])
            self.objset = ctypes.c_void_p()
            _lib.check(lib.gqp_objset_create(ctypes.cast(arr, ctypes.c_void_p), len(self.meshes), ctypes.byref(self.objset)))
        sp = torch.as_tensor(np.asarray(surface_points, np.float32), device=self.device)
        self.surface_points_per_object = sp  # (O, P, 3)
        self.surface_points_tensor = sp.repeat_interleave(self.batch_size_each, dim=0)  # :181-184
        n = len(self.meshes) * self.batch_size_each
        self.object_scale_tensor = torch.full((len(self.meshes), self.batch_size_each), self.object_scale, device=self.device)
        assert self.surface_points_tensor.shape[0] == n

    @property
    def cog(self):
        """object_model.py:64-68."""
        return self.surface_points_tensor.mean(1)

    def cal_distance(self, x, with_closest_points=False):
        """object_model.py:186-255 -> (distance (Bt,n) positive inside, outward normals (Bt,n,3))."""
        return ops.ObjectDistanceFn.apply(x, self)

    def __del__(self):
        try:
            lib = _lib.lib()
            if self.objset:
                lib.gqp_objset_destroy(self.objset)
            for m in self.meshes:
                lib.gqp_mesh_destroy(m)
            self.meshes, self.objset = [], None
        except Exception:
            pass
