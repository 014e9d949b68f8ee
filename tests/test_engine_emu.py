"""The whole iteration (FitEngine over the C-ABI, kernels run by the host emulator) vs the torch oracle of
scripts/fit.py:399-458 with injected noise — CPU only, tiny sizes."""
import numpy as np
import pytest
import torch

from graspqp_b200 import engine, hand_pack, synth
from oracle.pipeline import FitLoopOracle
import tol


def make_case(B=4, n_pts=64, n_c=12, seed=3):
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("sphere", 200, seed)
    fv = synth.face_vertices(v, f)
    sp = synth.sample_surface(v, f, n_pts, seed=seed)
    pose, idx = synth.random_grasp_init(pack, [v], B, n_c, seed)
    pose[:, :3] *= 0.7  # closer to the object: penetration and contact terms are active
    return pack, fv, sp, pose, idx


def noise(T, B, n_c, n_cand, seed=0):
    rng = np.random.default_rng(seed)
    return (rng.random((T, B, n_c), dtype=np.float32), rng.integers(0, n_cand, (T, B, n_c)).astype(np.int32),
            rng.random((T, B), dtype=np.float32))


def test_iteration_vs_oracle(emu):
    B, n_c, T = 4, 12, 3
    pack, fv, sp, pose, idx = make_case(B=B, n_c=n_c)
    eng = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    orc = FitLoopOracle(pack, [fv], [sp], B, pose, idx, dtype=torch.float64)
    np.testing.assert_allclose(eng.energy, orc.energy.numpy(), rtol=1e-4)
    for k, name in enumerate(engine.TERMS):
        np.testing.assert_allclose(eng.losses[k], orc.losses[name].numpy(), rtol=1e-4, atol=1e-6, err_msg=name)
    tol.assert_grad(eng.first_grad, orc.first_grad.numpy(), "emu first gradient")
    us, ir, ua = noise(T, B, n_c, len(pack["cand"]))
    for t in range(T):
        eng.step(us[t], ir[t], ua[t])
        out = orc.iterate(torch.tensor(us[t], dtype=torch.float64), torch.tensor(ir[t]).long(),
                          torch.tensor(ua[t], dtype=torch.float64))
        np.testing.assert_array_equal(eng.idx_new, out["idx_prop"].numpy())
        np.testing.assert_allclose(eng.pose_new, out["proposal"].numpy(), rtol=1e-4, atol=1e-5)
        tol.assert_energy(eng.new_energy, out["new_energy"].numpy(), f"emu step {t} new energy")
        tol.assert_grad(eng.grad_new, out["new_grad"].numpy(), f"emu step {t} gradient")
        np.testing.assert_allclose(eng.z, out["z"].numpy(), rtol=1e-3, atol=1e-4)
        np.testing.assert_array_equal(eng.accept.astype(bool), out["accept"].numpy())
        np.testing.assert_allclose(eng.temperature, out["temperature"].numpy(), rtol=1e-4)
        np.testing.assert_allclose(eng.energy, orc.energy.numpy(), rtol=tol.ENERGY_RTOL)
        np.testing.assert_allclose(eng.pose, orc.hand.hand_pose.detach().numpy(), rtol=1e-4, atol=1e-5)
    eng.close()


def test_short_trajectory_teacher_forced(emu):
    """the 500-step GPU trajectory check (tests/cases.py:check_trajectory) at emulator size"""
    import cases
    n_acc, n_tie, worst = cases.check_trajectory(emu, engine.NumpyBackend(), T=3, B=3, n_pts=48)
    assert n_tie == 0 and worst < tol.ENERGY_RTOL


def test_snapshot_restore_replays_the_same_iterations(emu):
    """restore(snapshot()) brings back the optimiser state: the Philox noise is keyed by the iteration number, so the
    replayed iterations are identical (bench.py times the same iterations twice this way)."""
    B, n_c = 3, 4
    pack, fv, sp, pose, idx = make_case(B=B, n_c=n_c, n_pts=64)
    eng = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], B, pose, idx.astype(np.int32))
    eng.step()
    snap = eng.snapshot()
    for _ in range(2):
        eng.step()
    want = (eng.energy.copy(), eng.pose.copy(), eng.idx.copy(), eng.iteration)
    eng.restore(snap)
    for _ in range(2):
        eng.step()
    np.testing.assert_array_equal(eng.energy, want[0])
    np.testing.assert_array_equal(eng.pose, want[1])
    np.testing.assert_array_equal(eng.idx, want[2])
    assert eng.iteration == want[3]


def test_sub_batch_check_with_batch_wide_statistics(emu):
    """logic of tests/cases.py:check_benched_state (the GPU test of BASELINE.json's configs at their benched shapes) at
    emulator size: a subset of the candidates vs the oracle seeded with the full batch's (g^2).mean(0), z-scores and
    QP exit iteration"""
    import cases
    import graspqp_b200.workload as wl
    wl.CONFIGS["tiny"] = ("allegro", ["sphere", "box"], 300, 4, 12)
    real = wl.build
    try:
        wl.build = lambda name, **kw: real(name, n_surface=48, **{k: v for k, v in kw.items() if v is not None})
        rep = cases.check_benched_state(emu, engine.NumpyBackend(), "tiny", iters=(0, 1), per_object=1)
    finally:
        wl.build = real
        del wl.CONFIGS["tiny"]
    assert all(r["energy_max"] <= 1e-4 and r["grad_max"] <= 1e-4 for r in rep.values())


def test_bounded_survivor_queue_overflow_is_detected(emu):
    """queue_per_candidate bounds the scratch of the penetration stage; a queue that fills up must be reported, never
    silently truncated"""
    from graspqp_b200 import _lib
    pack, fv, sp, pose, idx = make_case(B=4, n_c=4, n_pts=64)
    pose = pose.copy()
    pose[:, :3] *= 0.3  # hands deep inside the object: many samples survive the culling
    ref = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], 4, pose, idx.astype(np.int32))
    q, _, over = ref.pen_counters()
    assert q > 8 and not over
    ref.assert_ok()
    roomy = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], 4, pose, idx.astype(np.int32),
                             queue_per_candidate=64)
    np.testing.assert_array_equal(roomy.losses, ref.losses)
    roomy.assert_ok()
    tight = engine.FitEngine(emu, engine.NumpyBackend(), pack, [fv], sp[None], 4, pose, idx.astype(np.int32),
                             queue_per_candidate=1)
    assert tight.pen_counters()[2] == 1
    with pytest.raises(_lib.GqpError):
        tight.assert_ok()
    for e in (ref, roomy, tight):
        e.close()


@pytest.mark.parametrize("hand,n_c,w_wall", [("allegro", 12, 50.0), ("panda", 4, 0.0)])
def test_optional_terms_prior_and_wall(emu, hand, n_c, w_wall):
    """(f3) E_prior / E_wall (core/energy.py:68-77) through gqp_prior_wall vs the fp64 oracle; Panda's grasp axis is its
    forward axis (no grasp_axis passed, hand_model.py:470)"""
    import cases
    ep, ew = cases.check_optional_terms(emu, engine.NumpyBackend(), hand=hand, n_c=n_c, w_wall=w_wall)
    assert ep > 0 and (ew > 0) == (w_wall > 0)


def test_new_entry_points_check_their_arguments(emu):
    """(f3)/(f4) entry points: empty batches are a no-op, bad arguments are refused with a message (no launch)"""
    import ctypes
    from graspqp_b200 import _lib
    f, i = np.zeros(64, np.float32), np.zeros(64, np.int32)
    p = lambda a: a.ctypes.data
    S = _lib.GqpSpanParams(kind=_lib.SPAN_DEXGRASP, friction=0.2, obb_length=0.2, scale=100.0)
    assert emu.gqp_span_energy(ctypes.byref(S), None, None, None, None, None, None, None, 0, 0, 4, 4, None, None, None, None, None) == 0
    assert emu.gqp_span_energy(ctypes.byref(S), p(f), p(f), p(f), p(f), None, p(f), None, 0, 1, 1, 0, p(f), p(f), None, None, None) != 0
    assert b"n_c" in emu.gqp_last_error()
    M = _lib.GqpMalaParams(switch_possibility=0.5, starting_temperature=18.0, temperature_decay=0.95, step_size=0.005, mu=0.98,
                           annealing_period=30, stepsize_period=50, clip_grad=0)
    # an empty batch is fine; ema_in == ema_out is refused (every row must read the OLD accumulator)
    assert emu.gqp_anneal_propose(ctypes.byref(M), None, None, None, None, None, 0, None, 0, 25, 4, 92, None, None, 1, 1, None,
                                  None, None, None) == 0
    assert emu.gqp_anneal_propose(ctypes.byref(M), p(f), p(f), p(f), p(f), p(f), 0, p(i), 1, 25, 4, 92, None, None, 1, 1, p(f),
                                  p(i), None, None) != 0
    assert b"different buffers" in emu.gqp_last_error()
    assert emu.gqp_anneal_propose(ctypes.byref(M), p(f), p(f), p(f), p(f), p(f[32:]), -1, p(i), 1, 25, 4, 92, None, None, 1, 1,
                                  p(f), p(i), None, None) != 0  # negative step counter
    assert emu.gqp_anneal_accept(ctypes.byref(M), 0, 25, 4, 5, None, None, None, 1, 1, 0, None, None, None, None, None, None,
                                 None, None, None, None, None, 0, None) == 0
    assert emu.gqp_entropy_stats(p(f), 0, 16, p(f), p(f), p(f), p(f), None) != 0   # no rows: nothing to histogram
    assert emu.gqp_entropy_stats(p(f), 2, 16, p(f), p(f), None, p(f), None) != 0   # scratch is required
    assert emu.gqp_prior_wall(None, p(f), p(f), 1, None, None, None, 0, 1.0, 1.0, None, None, None, None, None) != 0
