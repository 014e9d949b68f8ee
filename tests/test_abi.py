"""The C-ABI boundary without a GPU: libgraspqp_b200.so loads and exports every symbol include/graspqp_b200.h
declares (no compute calls), the header and the ctypes table agree, and the argument / limit checks of the entry points
(host code shared by the CUDA build and the emulator build) return the documented error codes."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, ptr
from graspqp_b200 import _lib, build, hand_pack


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "graspqp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gqp_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_table_agree():
    assert _header_symbols() == sorted(_lib.SIGNATURES)


def test_cuda_library_exports_every_declared_symbol():
    path = build.build_cuda()  # nvcc cross-compiles for sm_100a without a GPU
    cdll = ctypes.CDLL(path)
    _lib.bind(cdll, strict=True)  # AttributeError if a declared symbol is missing
    assert cdll.gqp_abi_version() >= 1
    assert cdll.gqp_fc_workspace_bytes(8, 12, 12) > 8 * 12 * 3 * 48 * 8
    assert cdll.gqp_hand_distance_scratch_floats(2, 4, 100, 0) > 2 * 4 * 100


def test_package_refuses_to_compute_without_cuda():
    import torch
    from graspqp_b200 import engine
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="CUDA"):
            engine.TorchBackend("cpu")


def test_argument_and_limit_errors(emu):
    E_ARG, E_LIMIT = -1, -3
    f32 = lambda *s: np.zeros(s, np.float32)
    # meshes / object sets
    m = ctypes.c_void_p()
    assert emu.gqp_mesh_create(ptr(f32(9)), 0, ctypes.byref(m)) == E_ARG and b"F > 0" in emu.gqp_last_error()
    assert emu.gqp_mesh_create(None, 4, ctypes.byref(m)) == E_ARG
    s = ctypes.c_void_p()
    arr = (ctypes.c_void_p * 1)(None)
    assert emu.gqp_objset_create(ctypes.cast(arr, ctypes.c_void_p), 1, ctypes.byref(s)) == E_ARG
    assert b"mesh 0 is null" in emu.gqp_last_error()
    assert emu.gqp_object_distance_forward(None, 1, 1, None, 1.0, None, None, None, None, None) == E_ARG
    # QP: contact count limits, empty batches are fine
    P = _lib.GqpFcParams(friction=0.2, torque_weight=5.0, lower=1.0, upper=21.0, ridge=1e-4, svd_gain=0.1,
                         values_gain=2.0, values_offset=1e-2, eps=5e-2, max_iter=12, w_fc=1.0, w_dis=100.0)
    ws = np.zeros(1024, np.float64)
    assert emu.gqp_fc_solve(ctypes.byref(P), ptr(f32(3)), ptr(f32(3)), ptr(f32(3)), 1, 1, 17, ptr(ws), None) == E_LIMIT
    assert b"[2,16]" in emu.gqp_last_error()
    assert emu.gqp_fc_solve(ctypes.byref(P), ptr(f32(3)), ptr(f32(3)), ptr(f32(3)), 1, 1, 1, ptr(ws), None) == E_LIMIT
    assert emu.gqp_fc_solve(ctypes.byref(P), None, None, None, 0, 8, 12, None, None) == 0  # no candidates: nothing to do
    assert emu.gqp_fc_solve(ctypes.byref(P), ptr(f32(36)), ptr(f32(36)), ptr(f32(3)), 1, 1, 12, None, None) == E_ARG
    P.max_iter = 0
    assert emu.gqp_fc_solve(ctypes.byref(P), ptr(f32(36)), ptr(f32(36)), ptr(f32(3)), 1, 1, 12, ptr(ws), None) == E_ARG
    P.max_iter = 12
    assert emu.gqp_fc_finish(ctypes.byref(P), ptr(f32(36)), ptr(f32(36)), ptr(f32(12)), ptr(f32(36)), None, ptr(f32(3)),
                             1, 1, 12, ptr(ws), 1, None, None, None, None, None, None, None) == E_ARG  # no output asked
    # hand: table limits
    pack = dict(hand_pack.load_pack("allegro"))
    d, keep = hand_pack.make_desc(pack)
    d.n_frames = 1000
    h = ctypes.c_void_p()
    assert emu.gqp_hand_create(ctypes.byref(d), ctypes.byref(h)) != 0
    assert emu.gqp_hand_create(None, ctypes.byref(h)) == E_ARG
    # hand distance: scratch is required, empty batch is fine
    hand = hand_pack.create_hand(emu, pack)
    L = len(pack["mesh_frame"])
    assert emu.gqp_hand_distance(hand, ptr(f32(L * 12)), ptr(f32(3)), 1, 1, 1, 0, None, ptr(f32(1)), None, 1.0, None,
                                 None) == E_ARG
    assert b"scratch" in emu.gqp_last_error()
    assert emu.gqp_hand_distance(hand, None, None, 0, 5, 10, 0, None, None, None, 1.0, None, None) == 0
    assert emu.gqp_fk_forward(hand, None, 0, None, None) == 0
    emu.gqp_hand_destroy(hand)
