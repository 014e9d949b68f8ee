"""(f1) initialize_convex_hull on the device (csrc/gqp_init.cu) vs its restatement (oracle/init.py) under injected noise;
the truncated normal against torch.nn.init.trunc_normal_ itself (the one piece of initializations.py:15-193 whose
package is present in this image)."""
import numpy as np
import pytest
import torch

from graspqp_b200 import engine, hand_pack, initializations, synth
from oracle import init as oinit


def test_trunc_normal_matches_torch():
    g = torch.Generator().manual_seed(5)
    t = torch.empty(4000)
    torch.nn.init.trunc_normal_(t, 0.3, 0.2, -0.1 - 1e-6, 0.9 + 1e-6, generator=g)
    u = torch.rand(4000, generator=torch.Generator().manual_seed(5)).numpy()
    np.testing.assert_allclose(oinit.trunc_normal(u, 0.3, 0.2, -0.1 - 1e-6, 0.9 + 1e-6), t.numpy(), atol=3e-6)


def test_hull_is_outward_and_closed():
    v, f = synth.make_mesh("torus", 2000, 3)
    tris = oinit.convex_hull(v)
    n = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    assert ((n * (tris.mean(1) - v.mean(0))).sum(1) > 0).all()
    np.testing.assert_array_equal(tris, initializations.convex_hull(v))  # product copy of the same recipe


def check_device_init(lib, be, hand="allegro", B=16, n_c=12, pool_factor=20, seed=0):
    pack = hand_pack.load_pack(hand)
    hulls = [oinit.convex_hull(synth.make_mesh(k, 400, seed + i)[0]) for i, k in enumerate(("box", "hull"))]
    O, dof = len(hulls), len(pack["joints_lower"])
    rng = np.random.default_rng(seed)
    M = pool_factor * B
    u_pool = rng.random((O, M, 3), dtype=np.float32)
    u_pose = rng.random((O * B, 4 + dof), dtype=np.float32)
    idx_rand = rng.integers(0, len(pack["cand"]), (O * B, n_c)).astype(np.int32)
    want_pose, want_idx, picks = oinit.initialize_convex_hull(pack, hulls, B, n_c, u_pool, u_pose, idx_rand,
                                                              pool_factor=pool_factor)
    h = hand_pack.create_hand(lib, pack)
    di = initializations.DeviceInit(lib, be, pack, hulls, B, n_c, pool_factor=pool_factor)
    pose, idx, picked = be.empty((O * B, 9 + dof), np.float32), be.empty((O * B, n_c), np.int32), be.empty((O * B,), np.int32)
    di.run(h, pose, idx, be.from_host(u_pool), be.from_host(u_pose), be.from_host(idx_rand), picked=picked)
    np.testing.assert_array_equal(be.to_host(picked).reshape(O, B), np.stack(picks))  # FPS picks: bit-exact
    np.testing.assert_array_equal(be.to_host(idx), want_idx)
    got = be.to_host(pose)
    np.testing.assert_allclose(got[:, :3], want_pose[:, :3], atol=2e-7)   # translation
    np.testing.assert_allclose(got[:, 3:9], want_pose[:, 3:9], atol=2e-6)  # rotation columns
    np.testing.assert_allclose(got[:, 9:], want_pose[:, 9:], atol=3e-5)   # joints: two erfinv approximations, steep in the tails
    # sanity of the distribution: poses 5-10 cm (+1 cm inflation) off the hull, looking at it, joints inside their limits
    R0 = got[:, 3:6] / np.linalg.norm(got[:, 3:6], axis=1, keepdims=True)
    assert np.allclose(np.linalg.norm(got[:, 3:6], axis=1), 1, atol=1e-5) and np.abs((R0 * got[:, 6:9]).sum(1)).max() < 1e-5
    lo, hi = pack["joints_lower"], pack["joints_upper"]
    assert (got[:, 9:] >= lo - 2e-6).all() and (got[:, 9:] <= hi + 2e-6).all()
    # Philox path: deterministic in (seed, counter), different between calls
    p1, p2 = be.empty((O * B, 9 + dof), np.float32), be.empty((O * B, 9 + dof), np.float32)
    di.counter = 7
    di.run(h, p1, idx, seed=3)
    di.counter = 7
    di.run(h, p2, idx, seed=3)
    np.testing.assert_array_equal(be.to_host(p1), be.to_host(p2))
    di.run(h, p2, idx, seed=3)
    assert np.abs(be.to_host(p1) - be.to_host(p2)).max() > 1e-3
    assert np.isfinite(be.to_host(p2)).all()
    di.close()
    lib.gqp_hand_destroy(h)


@pytest.mark.parametrize("hand", ["allegro", "panda"])
def test_device_init_vs_oracle(emu, hand):
    check_device_init(emu, engine.NumpyBackend(), hand)


def test_engine_reset_step_with_device_init(emu):
    """FitEngine.reset_step: device init -> gqp_mala_reset -> evaluation -> forced accept of the masked rows"""
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("sphere", 200, 3)
    fv, sp = synth.face_vertices(v, f), synth.sample_surface(v, f, 48, seed=3)
    B, n_c = 4, 12
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, 3)
    eng = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    eng.enable_device_init([oinit.convex_hull(v)], pool_factor=8)
    eng.step()
    eng.reset_step(z_score_threshold=0.0)
    mask = eng.reset_mask.astype(bool)
    assert 0 < mask.sum() < B and int(eng.reset_count[0]) == mask.sum()
    np.testing.assert_array_equal(eng.pose[mask], eng.pose_init[mask])
    np.testing.assert_array_equal(eng.idx[mask], eng.idx_init[mask])
    assert (eng.step_count[mask] == 0).all() and (eng.step_count[~mask] == 2).all()
    assert (eng.ema[mask] == 0).all() and eng.accept[mask].all()
    eng.close()


@pytest.mark.skipif(not __import__("os").path.isdir("/root/reference"), reason="the reference is only present in the build container")
@pytest.mark.parametrize("hand", ["allegro", "panda"])
def test_oracle_init_matches_the_reference_function(hand):
    """The reference's OWN initialize_convex_hull (core/initializations.py:15-193), run as is, with its geometry back ends
    replaced by the oracle's pieces and its random draws injected:
        trimesh.Trimesh / .convex_hull / .face_normals / .nearest.on_surface -> hull triangles; closest point by the pinned
                                                                               point-triangle arithmetic (oracle.sdf)
        trimesh.sample.sample_surface_even                                   -> oracle.init.sample_pool on injected uniforms
        pytorch3d.ops.sample_farthest_points                                 -> oracle.init.fps
        transforms3d.euler.euler2mat                                         -> oracle.init.euler_rxyz
        torch.rand / randint / Tensor.uniform_ (inside trunc_normal_)        -> the injected draws
    What this pins is everything in between, which is the reference's code: the look-at construction, the inward normal from
    the closest hull point, distance / Euler jitter, the order of the rotations, the truncated normal (torch's own), the layout
    of hand_pose, and HandModel.set_parameters on the result."""
    import importlib
    import os
    import sys
    import types
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import ref_harness as rh
    from oracle import sdf as osdf
    rh.load_models()
    pack = hand_pack.load_pack(hand)
    B, n_c, pool_factor, seed = 8, 4, 100, 2  # the function hard-codes 100 * batch_size_each pool samples (:54)
    meshes = [synth.make_mesh(k, 400, seed + i) for i, k in enumerate(("box", "hull"))]
    hulls = [oinit.convex_hull(v) for v, _ in meshes]
    O, dof = len(hulls), len(pack["joints_lower"])
    rng = np.random.default_rng(seed)
    u_pool = rng.random((O, pool_factor * B, 3), dtype=np.float32)
    u_pose = rng.random((O * B, 4 + dof), dtype=np.float32)
    idx_rand = rng.integers(0, len(pack["cand"]), (O * B, n_c)).astype(np.int64)
    want_pose, want_idx, picks = oinit.initialize_convex_hull(pack, hulls, B, n_c, u_pool, u_pose, idx_rand, pool_factor=pool_factor)

    # ---- stand-ins for the absent geometry packages ----
    state = {"obj": -1}

    class FakeMesh:
        def __init__(self, vertices=None, faces=None, tris=None):
            if tris is None:
                tris = np.asarray(vertices, np.float64)[np.asarray(faces)]
            self.tris = np.asarray(tris, np.float32)
            self.vertices = self.tris.reshape(-1, 3).astype(np.float64)
            self.faces = np.arange(len(self.vertices)).reshape(-1, 3)
            self.face_normals = oinit.hull_tables(self.tris)[0].astype(np.float64)
            self.nearest = types.SimpleNamespace(on_surface=self._on_surface)

        @property
        def convex_hull(self):
            return self

        def remove_degenerate_faces(self):
            return np.ones(len(self.faces), bool)

        def _on_surface(self, pts):
            _, _, _, cl = osdf.compute_sdf_port(np.asarray(pts, np.float32), self.tris)
            return cl.astype(np.float64), None, None

    def sample_surface_even(mesh, count):
        state["obj"] += 1
        normals, cdf = oinit.hull_tables(mesh.tris)
        pts, face = oinit.sample_pool(mesh.tris, normals, cdf, u_pool[state["obj"]][:count], inflate=0.0)
        return pts.astype(np.float64), face

    def sample_farthest_points(points, K, random_start_point=False):
        assert not random_start_point
        p32 = points[0].numpy().astype(np.float32)
        sel = oinit.fps(p32, K)
        return torch.as_tensor(p32[sel]).unsqueeze(0), torch.as_tensor(sel).unsqueeze(0)

    def euler2mat(ai, aj, ak, axes="sxyz"):
        assert axes == "rxyz"
        return oinit.euler_rxyz(np.array([float(ai)], np.float32), np.array([float(aj)], np.float32),
                                np.array([float(ak)], np.float32))[0]

    tm = sys.modules["trimesh"]
    tm.Trimesh, tm.sample = FakeMesh, types.SimpleNamespace(sample_surface_even=sample_surface_even)
    sys.modules["pytorch3d.ops"].sample_farthest_points = sample_farthest_points
    t3d = sys.modules.setdefault("transforms3d", types.ModuleType("transforms3d"))
    t3d.euler = types.SimpleNamespace(euler2mat=euler2mat)
    mod = importlib.import_module("graspqp.core.initializations")

    # ---- injected noise: torch.rand (4 per object), torch.randint (contacts), Tensor.uniform_ (trunc_normal_, per joint) ----
    calls = {"rand": 0, "joint": 0}

    class Noise:
        def __getattr__(self, k):
            return getattr(torch, k)

        def rand(self, size, **kw):
            o, c = divmod(calls["rand"], 4)
            calls["rand"] += 1
            return torch.tensor(u_pose[o * B:(o + 1) * B, c])

        def randint(self, high, size, **kw):
            return torch.tensor(idx_rand)

    real_uniform = torch.Tensor.uniform_

    def uniform_(self, a=0.0, b=1.0, **kw):  # torch/nn/init.py: tensor.uniform_(2l - 1, 2u - 1) of column `joint`
        j = calls["joint"]
        calls["joint"] += 1
        with torch.no_grad():
            self.copy_(torch.tensor(u_pose[:, 4 + j]) * (b - a) + a)
        return self

    h = rh.reference_hand_model(pack)
    ax = {"x": (1, 0, 0), "y": (0, 1, 0), "z": (0, 0, 1)}
    h.forward_axis = torch.tensor(ax[str(pack["axes"][0])], dtype=torch.float)
    h.up_axis = torch.tensor(ax[str(pack["axes"][1])], dtype=torch.float)
    h.default_state = torch.tensor(np.asarray(pack["default_state"], np.float32))
    obj = types.SimpleNamespace(object_mesh_list=[FakeMesh(tris=t) for t in hulls], batch_size_each=B,
                                object_scale_tensor=torch.ones(O, B), data_root_path="/data/release")
    D = dict(oinit.DEFAULTS)
    args = types.SimpleNamespace(n_contact=n_c, **{k: D[k] for k in D if k not in ("inflate", "pool_factor")})
    mod.torch, torch.Tensor.uniform_ = Noise(), uniform_
    try:
        mod.initialize_convex_hull(h, obj, args)
    finally:
        mod.torch, torch.Tensor.uniform_ = torch, real_uniform
    got = h.hand_pose.detach().numpy()
    print(f"[init] {hand}: reference function vs oracle: translation {np.abs(got[:, :3] - want_pose[:, :3]).max():.1e}, "
          f"rotation {np.abs(got[:, 3:9] - want_pose[:, 3:9]).max():.1e}, joints {np.abs(got[:, 9:] - want_pose[:, 9:]).max():.1e}; "
          f"calls: rand {calls['rand']}, uniform_ {calls['joint']}, surface samplings {state['obj'] + 1}")
    assert calls["rand"] == 4 * O and calls["joint"] == dof and state["obj"] + 1 == O
    np.testing.assert_array_equal(h.contact_point_indices.numpy(), want_idx)
    np.testing.assert_allclose(got[:, :3], want_pose[:, :3], atol=1e-6)    # translation
    np.testing.assert_allclose(got[:, 3:9], want_pose[:, 3:9], atol=5e-6)  # two rotation columns
    np.testing.assert_allclose(got[:, 9:], want_pose[:, 9:], atol=3e-6)    # joints (torch's own erfinv on both sides here)
    assert h.contact_points.shape == (O * B, n_c, 3)                       # HandModel.set_parameters ran on the result
