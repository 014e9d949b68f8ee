"""K1 kernel text (csrc/gqp_sdf.cu) run by the host emulator vs the oracle (bit-exact) — CPU only."""
import numpy as np
import pytest

from conftest import ptr
from graspqp_b200 import synth
from oracle import sdf as osdf


def _run(emu, pts, fv):
    P, F = len(pts), len(fv)
    d2 = np.empty(P, np.float32); sg = np.empty(P, np.int32)
    nr = np.empty((P, 3), np.float32); cl = np.empty((P, 3), np.float32)
    rc = emu.gqp_sdf_forward(ptr(pts), ptr(fv), P, F, ptr(d2), ptr(sg), ptr(nr), ptr(cl), None)
    assert rc == 0, emu.gqp_last_error()
    return d2, sg, nr, cl


@pytest.mark.parametrize("kind,nf,P", [("sphere", 80, 300), ("box", 600, 257), ("torus", 1100, 64), ("hull", 3, 40)])
def test_sdf_emu_bit_exact(emu, oracle_built, kind, nf, P):
    if kind == "hull":  # F not a multiple of 4 and smaller than a tile: exercises the non-bulk tail
        v, f = synth.make_mesh("sphere", 40, 1)
        f = f[:nf]
    else:
        v, f = synth.make_mesh(kind, nf, 2)
    fv = synth.face_vertices(v, f)
    rng = np.random.default_rng(0)
    pts = (rng.uniform(-0.15, 0.15, (P, 3))).astype(np.float32)
    pts[: P // 4] = (fv.reshape(-1, 3)[rng.integers(0, len(fv) * 3, P // 4)] * 1.0001).astype(np.float32)
    got = _run(emu, pts, fv)
    want = osdf.compute_sdf_port(pts, fv)
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g, w)
    if osdf.have_ref():
        for g, w in zip(got, osdf.compute_sdf_ref(pts, fv)):
            np.testing.assert_array_equal(g, w)


def test_sdf_emu_unaligned_and_errors(emu, oracle_built):
    v, f = synth.make_mesh("sphere", 80, 1)
    fv = synth.face_vertices(v, f)
    buf = np.empty(fv.size + 1, np.float32)
    fv_un = buf[1:].reshape(fv.shape)  # 4-byte aligned only -> plain-load staging path
    fv_un[...] = fv
    pts = np.random.default_rng(1).uniform(-0.1, 0.1, (33, 3)).astype(np.float32)
    got = _run(emu, pts, fv_un)
    for g, w in zip(got, osdf.compute_sdf_port(pts, fv)):
        np.testing.assert_array_equal(g, w)
    assert emu.gqp_sdf_forward(ptr(pts), ptr(fv), 33, 0, None, None, None, None, None) == -1
    assert b"F > 0" in emu.gqp_last_error()
    assert emu.gqp_sdf_forward(ptr(pts), ptr(fv), 0, 5, None, None, None, None, None) == 0


def test_sdf_backward_emu(emu):
    rng = np.random.default_rng(2)
    p, c = rng.normal(size=(50, 3)).astype(np.float32), rng.normal(size=(50, 3)).astype(np.float32)
    g = rng.normal(size=50).astype(np.float32)
    out = np.empty((50, 3), np.float32)
    assert emu.gqp_sdf_backward(ptr(g), ptr(p), ptr(c), 50, ptr(out), None) == 0
    np.testing.assert_array_equal(out, (p - c) * (2 * g)[:, None])


@pytest.mark.parametrize("kind,nf,P", [("sphere", 3, 5), ("sphere", 8, 1), ("sphere", 9, 37), ("box", 70, 33),
                                        ("torus", 1100, 41), ("hull", 3000, 24)])
def test_mesh_sdf_gang_emu_bit_exact(emu, oracle_built, kind, nf, P):
    """gqp_mesh_sdf_forward (8-lane gangs over the 8-wide BVH) == brute force, winner face included; covers a root that
    is a leaf, a single partial wide node, ragged last blocks."""
    import cases
    from graspqp_b200 import engine
    v, f = synth.make_mesh(kind if nf > 9 else "sphere", max(nf, 40), 3)
    fv = synth.face_vertices(v, f)[:nf] if nf <= 9 else synth.face_vertices(v, f)
    pts = cases.query_points(fv, P, seed=nf)
    want = osdf.compute_sdf_port(pts, fv, with_face=True)
    got = cases.mesh_sdf_forward(emu, engine.NumpyBackend(), pts, fv)
    for g, w in zip(got, want):
        np.testing.assert_array_equal(g, w)
