"""Drop-in for the `torchsdf` package (thirdparty/TorchSDF/torchsdf/sdf.py): same two functions, same return contract."""
import torch

from . import ops


def index_vertices_by_faces(vertices_features, faces):
    """sdf.py:5-31: (V,3) vertices + (F,3) faces -> (F,3,3) face vertices."""
    assert vertices_features.ndim == 2 and faces.ndim == 2
    return vertices_features[faces.long()]


def compute_sdf(pointclouds, face_vertices):
    """sdf.py:34-68: -> (squared distance [differentiable], sign int32 (+1 outside / -1 inside by the nearest face's
    normal), unit(p - closest), closest point).  CPU inputs are moved to CUDA and back like the reference does."""
    was_cpu = not pointclouds.is_cuda
    if was_cpu:
        pointclouds, face_vertices = pointclouds.cuda(), face_vertices.cuda()
    out = ops.SdfFn.apply(pointclouds.reshape(-1, 3), face_vertices.reshape(-1, 3, 3))
    return tuple(o.cpu() for o in out) if was_cpu else out
