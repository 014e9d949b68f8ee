"""ctypes binding of libgraspqp_b200.so (include/graspqp_b200.h).  No CPU fallback: if the library is missing or fails to
load, importing anything that computes raises."""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_uint8, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgraspqp_b200.so")


class GqpHandDesc(ctypes.Structure):
    _fields_ = [(n, c_int32) for n in ("n_frames", "n_dofs", "n_links", "n_cand", "n_spheres")] + [
        (n, c_void_p) for n in (
            "frame_parent", "frame_jtype", "frame_dof", "frame_axis", "frame_joff", "frame_loff", "link_frame",
            "face_offset", "face_verts", "cand", "cand_normal", "cand_link", "spheres", "sphere_link", "joints_lower",
            "joints_upper")] + [("mlp_hidden", c_int32), ("mlp_out", c_int32), ("mlp_lo", c_float),
                                ("mlp_hi", c_float)] + [(n, c_void_p) for n in ("mlp_w1", "mlp_b1", "mlp_w2", "mlp_b2", "frame_gain")]


class GqpFcParams(ctypes.Structure):
    _fields_ = [(n, c_float) for n in ("friction", "torque_weight", "lower", "upper", "ridge", "svd_gain",
                                       "values_gain", "values_offset", "eps")] + [("max_iter", c_int32),
                                                                                 ("w_fc", c_float), ("w_dis", c_float)]


class GqpMalaParams(ctypes.Structure):
    _fields_ = [(n, c_float) for n in ("switch_possibility", "starting_temperature", "temperature_decay", "step_size",
                                       "mu")] + [(n, c_int32) for n in ("annealing_period", "stepsize_period",
                                                                        "clip_grad")]


class GqpSpanParams(ctypes.Structure):
    _fields_ = [("kind", c_int32), ("torque_weight", c_float), ("friction", c_float), ("obb_length", c_float),
                ("scale", c_float), ("enable_density", c_int32), ("w_fc", c_float), ("w_dis", c_float)]


SPAN_DEXGRASP, SPAN_TDG = 1, 2


class GqpInitParams(ctypes.Structure):
    _fields_ = [(n, c_float) for n in ("distance_lower", "distance_upper", "rotate_lower", "rotate_upper", "pitch_lower",
                                       "pitch_upper", "tilt_lower", "tilt_upper", "jitter_strength", "inflate")] + [
        ("pool_factor", c_int32), ("forward_axis", c_float * 3), ("up_axis", c_float * 3), ("default_state", c_float * 32)]


P = c_void_p  # device / host pointers are passed as integers (tensor.data_ptr())

# name -> (restype, argtypes); every symbol include/graspqp_b200.h declares
SIGNATURES = {
    "gqp_abi_version": (c_int, []),
    "gqp_last_error": (c_char_p, []),
    "gqp_launch_count": (c_uint64, []),
    "gqp_sdf_forward": (c_int, [P, P, c_int, c_int, P, P, P, P, P]),
    "gqp_sdf_backward": (c_int, [P, P, P, c_int, P, P]),
    "gqp_mesh_create": (c_int, [P, c_int, POINTER(c_void_p)]),
    "gqp_mesh_destroy": (None, [c_void_p]),
    "gqp_mesh_num_faces": (c_int, [c_void_p]),
    "gqp_mesh_sdf_forward": (c_int, [c_void_p, P, c_int, P, P, P, P, P, P]),
    "gqp_objset_create": (c_int, [P, c_int, POINTER(c_void_p)]),
    "gqp_objset_destroy": (None, [c_void_p]),
    "gqp_object_distance_forward": (c_int, [c_void_p, c_int, c_int, P, c_float, P, P, P, P, P]),
    "gqp_object_distance_backward": (c_int, [P, P, P, P, c_int, c_float, P, P]),
    "gqp_hand_create": (c_int, [POINTER(GqpHandDesc), POINTER(c_void_p)]),
    "gqp_hand_destroy": (None, [c_void_p]),
    "gqp_fk_forward": (c_int, [c_void_p, P, c_int, P, P]),
    "gqp_fk_backward": (c_int, [c_void_p, P, P, P, c_int, P, P]),
    "gqp_contacts_forward": (c_int, [c_void_p, P, P, c_int, c_int, P, P, P]),
    "gqp_contacts_backward": (c_int, [c_void_p, P, P, P, c_int, c_int, P, P]),
    "gqp_fk_contacts_forward": (c_int, [c_void_p, P, P, c_int, c_int, P, P, P, P]),
    "gqp_fk_contacts_backward": (c_int, [c_void_p, P, P, P, P, P, P, c_int, c_int, P, P]),
    "gqp_hand_distance": (c_int, [c_void_p, P, P, c_int, c_int, c_int, c_int, P, P, P, c_float, P, P]),
    "gqp_hand_distance_full": (c_int, [c_void_p, P, P, c_int, c_int, c_int, c_int, P, P, P, P, P]),
    "gqp_hand_distance_backward": (c_int, [c_void_p, P, P, c_int, c_int, c_int, c_int, P, P, P, P, P]),
    "gqp_hand_distance_scratch_floats": (c_uint64, [c_int, c_int, c_int, c_int]),
    "gqp_hand_distance_scratch_floats_bounded": (c_uint64, [c_int, c_int, c_int, c_int, c_int]),
    "gqp_hand_distance_bounded": (c_int, [c_void_p, P, P, c_int, c_int, c_int, c_int, P, P, c_float, P, c_int, P]),
    "gqp_hand_distance_counter_offset": (c_uint64, [c_int, c_int, c_int, c_int, P]),
    "gqp_hand_pen_info": (c_int, [c_void_p, P]),
    "gqp_hand_link_grid": (c_int, [c_void_p, c_int, P, P, P, c_int]),
    "gqp_hand_link_planes": (c_int, [c_void_p, c_int, P, c_int, P]),
    "gqp_self_pen_joints": (c_int, [c_void_p, P, P, c_int, P, P, c_float, c_float, P, P, P]),
    "gqp_prior_wall": (c_int, [c_void_p, P, P, c_int, POINTER(c_float * 3), P, P, c_int, c_float, c_float, P, P, P, P, P]),
    "gqp_fc_dis_energy": (c_int, [POINTER(GqpFcParams), P, P, P, P, P, P, c_int, c_int, c_int, P, P, P, P, P, P, P, P]),
    "gqp_fc_stop_iter": (c_int, [P, c_int, c_int, c_float, P, P]),
    "gqp_fc_workspace_bytes": (c_uint64, [c_int, c_int, c_int]),
    "gqp_fc_solve": (c_int, [POINTER(GqpFcParams), P, P, P, c_int, c_int, c_int, P, P]),
    "gqp_fc_finish": (c_int, [POINTER(GqpFcParams), P, P, P, P, P, P, c_int, c_int, c_int, P, c_int, P, P, P, P, P, P, P]),
    "gqp_span_energy": (c_int, [POINTER(GqpSpanParams), P, P, P, P, P, P, P, c_int, c_int, c_int, c_int, P, P, P, P, P]),
    "gqp_mala_stats": (c_int, [P, P, c_int, c_int, c_int, c_int, P, P, P, P]),
    "gqp_mala_stats_noalloc": (c_int, [P, P, c_int, c_int, c_int, c_int, P, P, P]),
    "gqp_mala_stats_noalloc_max_floats": (c_uint64, []),
    "gqp_mala_propose": (c_int, [POINTER(GqpMalaParams), P, P, P, P, P, P, c_int, c_int, c_int, c_int, P, P, c_uint64,
                                 c_uint64, P, P, P, P]),
    "gqp_mala_accept": (c_int, [POINTER(GqpMalaParams), c_int, c_int, c_int, c_int, P, P, P, P, P, c_uint64, c_uint64,
                                P, P, P, P, P, P, P, P, P, P, P, P, c_int, P]),
    "gqp_anneal_propose": (c_int, [POINTER(GqpMalaParams), P, P, P, P, P, c_int, P, c_int, c_int, c_int, c_int, P, P,
                                   c_uint64, c_uint64, P, P, P, P]),
    "gqp_anneal_accept": (c_int, [POINTER(GqpMalaParams), c_int, c_int, c_int, c_int, P, P, P, c_uint64, c_uint64, c_int,
                                  P, P, P, P, P, P, P, P, P, P, P, c_int, P]),
    "gqp_entropy_stats": (c_int, [P, c_int, c_int, P, P, P, P, P]),
    "gqp_mala_reset": (c_int, [c_int, c_int, c_int, P, c_float, P, P, P, P, P, P, P, P, P, P, P]),
    "gqp_init_create": (c_int, [P, P, c_int, POINTER(c_void_p)]),
    "gqp_init_destroy": (None, [c_void_p]),
    "gqp_init_scratch_bytes": (c_uint64, [c_int, c_int, c_int]),
    "gqp_init_convex_hull": (c_int, [c_void_p, c_void_p, POINTER(GqpInitParams), c_int, c_int, P, P, P, c_uint64,
                                     c_uint64, P, P, P, P, P]),
    "gqp_memset_zero": (c_int, [P, c_uint64, P]),
    "gqp_fma_peak_probe": (c_int, [c_int, P, POINTER(c_uint64), P]),
    "gqp_dfma_peak_probe": (c_int, [c_int, P, POINTER(c_uint64), P]),
}


def bind(cdll, strict=True):
    """Attach prototypes; raises AttributeError if a declared symbol is not exported (strict)."""
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(cdll, name)
        except AttributeError:
            if strict:
                raise
            continue
        fn.restype = res
        fn.argtypes = args
    return cdll


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python graspqp_b200/build.py cuda` "
                "(graspqp_b200 has no CPU or PyTorch fallback)")
        _lib = bind(ctypes.CDLL(LIB_PATH))
    return _lib


class GqpError(RuntimeError):
    pass


def check(rc, cdll=None):
    if rc != 0:
        msg = (cdll or lib()).gqp_last_error()
        raise GqpError(f"graspqp_b200 error {rc}: {msg.decode() if msg else ''}")
