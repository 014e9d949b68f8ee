"""Parity tests proper: the CUDA path through the C-ABI (libgraspqp_b200.so) vs the oracle.  Run with -m gpu on a B200."""
import ctypes

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu(oracle_built):
    from graspqp_b200 import _lib, engine
    assert torch.cuda.is_available()
    return _lib.lib(), engine.TorchBackend("cuda:0")


@pytest.mark.parametrize("kind,nf,P", [("sphere", 80, 1000), ("box", 2000, 20000), ("torus", 10000, 30011),
                                       ("hull", 10000, 4096), ("cylinder", 2051, 777)])
def test_sdf_bit_exact(gpu, kind, nf, P):
    cases.check_sdf_bit_exact(*gpu, kind, nf, P)


def test_sdf_self_consistency_full_size(gpu):
    """thirdparty/TorchSDF/tests/normal.py:36-39 law at 50k faces x 200k points: normal * 2 sqrt(d2) == backward(1);
    BVH and brute force agree bit for bit (winner included)."""
    lib, be = gpu
    from graspqp_b200 import synth
    v, f = synth.make_mesh("hull", 50000, 7)
    fv = synth.face_vertices(v, f)
    pts = cases.query_points(fv, 200000, 1)
    d2, sg, nr, cl = cases.sdf_forward(lib, be, pts, fv)
    d2b, sgb, nrb, clb, fi = cases.mesh_sdf_forward(lib, be, pts, fv)
    np.testing.assert_array_equal(d2, d2b); np.testing.assert_array_equal(sg, sgb)
    np.testing.assert_array_equal(cl, clb); np.testing.assert_array_equal(nr, nrb)
    np.testing.assert_allclose(nr * 2 * np.sqrt(d2)[:, None], 2 * (pts - cl), atol=5e-7)
    np.testing.assert_allclose(((pts - cl) ** 2).sum(-1), d2, rtol=1e-5, atol=1e-12)
    assert set(np.unique(sg)) <= {-1, 1}


def test_hand_distance(gpu):
    cases.check_hand_distance(*gpu, B=448)


@pytest.mark.parametrize("hand", ["shadow_hand", "panda", "robotiq2"])
def test_hand_distance_other_hands(gpu, hand):
    """Culled E_pen == un-culled (every sample searched) for 448 poses x 2500 samples = 1.1 M (pose, sample) pairs per
    hand, samples all around the links: the Panda fingers carry a concave crease sharper than 90 degrees that the
    far-field culling proof (gqp_mesh_is_nice) has to get right."""
    cases.check_hand_distance(*gpu, B=448, hand=hand)


@pytest.mark.parametrize("hand", ["allegro", "shadow_hand", "panda", "robotiq2"])
def test_culling_proof_outside_cells_have_positive_sign(gpu, hand):
    """Direct test of the proof behind the culling: 16 random points + 8 corners in EVERY set cell of every link's outside
    grid and 400k points outside the link boxes -> the brute-force kernel (bit-exact TorchSDF arithmetic) says +1."""
    lib, be = gpu
    sign = lambda pts, fv: cases.sdf_forward(lib, be, pts, fv)[1]
    n_links, n_cells, n_points = cases.check_culling_proof(lib, sign, hand, per_cell=16, n_far=200000)
    assert n_links >= 1 and n_points > 100000


@pytest.mark.parametrize("name,iters,kw", [
    ("config2", (0, 50, 500), {}),
    ("config3", (0, 50, 300), {}),
    ("config4", (0, 50, 200), {"n_objects": 8}),
    ("config4_robotiq2", (0, 50, 200), {"n_objects": 8}),
    ("config5", (0, 50, 300), {}),
])
def test_parity_at_benched_shapes(gpu, name, iters, kw):
    """BASELINE.json's configs at the shapes bench.py runs (full candidate count per object, 2500 surface samples,
    10k / 50k-face meshes, batch-wide QP exit over the whole batch): a random subset of the candidates of the state
    after 0 / 50 / 300-500 iterations against the fp64 oracle seeded with the full batch's statistics."""
    rep = cases.check_benched_state(*gpu, name=name, iters=iters, **kw)
    assert all(r["energy_max"] <= 1e-4 and r["grad_max"] <= 1e-4 for r in rep.values())


def test_engine_vs_oracle(gpu):
    n_acc = cases.check_engine_vs_oracle(*gpu, B=8, n_pts=256, T=4)
    assert n_acc > 0


def test_engine_vs_oracle_hull_nc16(gpu):
    cases.check_engine_vs_oracle(*gpu, B=6, n_pts=200, n_c=16, T=3, seed=11, hull=True)


def test_philox_run_is_deterministic(gpu):
    lib, be = gpu
    from graspqp_b200 import engine, hand_pack, synth
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("box", 1000, 2)
    fv, sp = synth.face_vertices(v, f), synth.sample_surface(v, f, 512, seed=2)
    pose, idx = synth.random_grasp_init(pack, [v, v], 64, 12, 2)
    runs = []
    for _ in range(2):
        eng = engine.FitEngine(lib, be, pack, [fv, fv], np.stack([sp, sp]), 64, pose, idx.astype(np.int32), seed=5)
        for _ in range(10):
            eng.step()
        runs.append((be.to_host(eng.energy), be.to_host(eng.idx), be.to_host(eng.accept)))
        eng.close()
    np.testing.assert_array_equal(runs[0][1], runs[1][1])
    np.testing.assert_array_equal(runs[0][2], runs[1][2])
    np.testing.assert_allclose(runs[0][0], runs[1][0], rtol=1e-4)
    assert np.isfinite(runs[0][0]).all()


# ---- against the committed golden vectors: outputs of the reference's own code (tests/golden/make_golden.py) ----
def test_golden_sdf(gpu):
    import golden_cases as gc
    gc.check_sdf(*gpu)


def test_golden_fk(gpu):
    import golden_cases as gc
    gc.check_fk(*gpu)


def test_golden_metric(gpu):
    import golden_cases as gc
    gc.check_metric(*gpu)


def test_golden_fit_iterations(gpu):
    import golden_cases as gc
    assert gc.check_fit(*gpu, T=4) > 0


def test_golden_reset_iterations(gpu):
    """M3 (reset_envs, reset mask, forced accept), clip_grad and NaN gradients against the reference's own MalaStar"""
    import golden_cases as gc
    assert gc.check_fit_reset(*gpu) == 4


@pytest.mark.parametrize("hand,B", [("allegro", 256), ("shadow_hand", 128), ("robotiq2", 64)])
def test_device_init_vs_oracle(gpu, hand, B):
    """(f1) initialize_convex_hull on the device at the reference's pool factor (100 x B samples per object): FPS picks
    bit-exact, poses within 2e-6 of the restatement under injected noise"""
    import test_init
    test_init.check_device_init(*gpu, hand=hand, B=B, pool_factor=100)


@pytest.mark.parametrize("hand", ["allegro", "shadow_hand", "panda", "robotiq2"])
def test_export_jacobian_and_joint_velocities(gpu, hand):
    """(f2) contact Jacobian from the FK backward kernels == autograd through the restated HandModel; joint velocities ==
    the reference's damped pseudo-inverse formula in its own (hand) frame"""
    import test_export
    test_export.check_jacobian_and_velocities(*gpu, hand, B=16, n_c=12 if hand in ("allegro", "shadow_hand") else 4)


def test_trajectory_500_steps_teacher_forced(gpu):
    n_acc, n_tie, worst = cases.check_trajectory(*gpu, T=500)
    print(f"500-step trajectory: {n_acc} accepted, {n_tie} borderline accept mismatches, worst energy error {worst:.2e}")
    assert n_acc > 500


@pytest.mark.parametrize("hand,n_c", [("shadow_hand", 12), ("robotiq2", 4), ("panda", 4)])
def test_trajectory_other_hands(gpu, hand, n_c):
    """150 teacher-forced steps with the Shadow Hand (24 dof, 18 links), the Robotiq 2F-140 (coupling network in the FK
    backward) and the Panda gripper (both small-QP regime, n_c = 4): same per-step tolerances as the Allegro trajectory,
    E_fc on flagged conditioning events within 1e-5 of the oracle on the engine's own inputs.  (Round 1 needed 5e-3 at
    n_c = 4: the svd term was taken from a Cholesky factor of F F', which loses sigma_min below 1e-8 sigma_max on
    rank-deficient grasp matrices — tests/test_fc_emu.py::test_rank_deficient_grasp_matrix_small_qp.)"""
    n_acc, n_tie, worst = cases.check_trajectory(*gpu, T=150, n_c=n_c, hand=hand, fc_tol=1e-5)
    print(f"{hand}: {n_acc} accepted, {n_tie} borderline, worst energy error {worst:.2e}")
    assert n_acc > 50


@pytest.mark.parametrize("hand,n_c", [("shadow_hand", 12), ("panda", 4), ("robotiq2", 4)])
def test_engine_vs_oracle_other_hands(gpu, hand, n_c):
    cases.check_engine_vs_oracle(*gpu, B=6, n_pts=200, n_c=n_c, T=3, seed=5, hand=hand)
