"""TEST INFRASTRUCTURE. ctypes front-ends of oracle/_build/libsdf_oracle.so (our C restatement of TorchSDF) and of
oracle/_ref/libtorchsdf_ref.so (the reference kernel body itself, host-compiled).  Same return contract as
/root/reference/thirdparty/TorchSDF/torchsdf/sdf.py:34-68: (dist^2, sign int32, unit(p-c), closest point)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)


def _load(rel):
    path = os.path.join(_HERE, rel)
    if not os.path.exists(path):
        subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh")])
    return ctypes.CDLL(path)


_port = None
_ref = None


def _p(a, t):
    return a.ctypes.data_as(t)


def compute_sdf_port(points, face_verts, with_face=False):
    """Our restatement (oracle/sdf_oracle.c)."""
    global _port
    if _port is None:
        _port = _load("_build/libsdf_oracle.so")
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    fv = np.ascontiguousarray(face_verts, np.float32).reshape(-1, 9)
    P, F = len(pts), len(fv)
    d2 = np.empty(P, np.float32); sg = np.empty(P, np.int32)
    nr = np.empty((P, 3), np.float32); cl = np.empty((P, 3), np.float32); fi = np.empty(P, np.int32)
    _port.sdf_oracle_forward(_p(pts, _f32p), _p(fv, _f32p), P, F, _p(d2, _f32p), _p(sg, _i32p), _p(nr, _f32p),
                             _p(cl, _f32p), _p(fi, _i32p))
    return (d2, sg, nr, cl, fi) if with_face else (d2, sg, nr, cl)


def set_threads(n):
    """OpenMP team size of the C restatement (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    global _port
    if _port is None:
        _port = _load("_build/libsdf_oracle.so")
    _port.sdf_oracle_set_threads(int(n))
    return int(_port.sdf_oracle_get_threads())


def have_ref():
    return os.path.exists(os.path.join(_HERE, "_ref", "libtorchsdf_ref.so"))


def compute_sdf_ref(points, face_verts):
    """The reference's own kernel body, compiled for the host (oracle/build_ref.sh)."""
    global _ref
    if _ref is None:
        _ref = ctypes.CDLL(os.path.join(_HERE, "_ref", "libtorchsdf_ref.so"))
    pts = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    fv = np.ascontiguousarray(face_verts, np.float32).reshape(-1, 9)
    P, F = len(pts), len(fv)
    d2 = np.empty(P, np.float32); sg = np.empty(P, np.int32)
    nr = np.empty((P, 3), np.float32); cl = np.empty((P, 3), np.float32)
    _ref.torchsdf_ref_forward(_p(pts, _f32p), _p(fv, _f32p), P, F, _p(d2, _f32p), _p(sg, _i32p), _p(nr, _f32p),
                              _p(cl, _f32p))
    return d2, sg, nr, cl
