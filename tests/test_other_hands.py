"""Hands beyond Allegro (Shadow 24 dof / 18 mesh links, Panda 1 dof driving two prismatic fingers; 8(f3):
Robotiq 3F 11 dof, ability hand 6 parameters driving 10 joints through mimic gains):
the whole iteration through the emulated kernels vs the fp64 oracle.  FK / contact transforms of these hands are pinned
against pytorch_kinematics itself in tests/test_golden.py (fk_ref.npz)."""
import numpy as np
import pytest

import cases
from graspqp_b200 import engine


@pytest.mark.parametrize("hand,n_c", [("shadow_hand", 12), ("panda", 4), ("robotiq2", 4), ("robotiq3", 12),
                                      ("ability_hand", 8)])
def test_iteration_vs_oracle_other_hands(emu, hand, n_c):
    n_acc = cases.check_engine_vs_oracle(emu, engine.NumpyBackend(), B=3, n_pts=48, n_c=n_c, T=2, seed=5, hand=hand)
    assert n_acc >= 0


@pytest.mark.parametrize("hand", ["shadow_hand", "panda", "robotiq2", "robotiq3", "ability_hand"])
def test_culled_pen_equals_unculled_other_hands(emu, hand):
    """H4 on poses pushed into the object: the culled path (outside grids, link boxes) must give the E_pen of the path
    that searches every (sample, link) pair.  Guards the culling proofs of gqp_bvh.cpp (concave creases of the Panda
    fingers, nested shells of the Shadow thumb base); the full-size version runs on the GPU."""
    cases.check_hand_distance(emu, engine.NumpyBackend(), B=4, n_pts=64, hand=hand)
