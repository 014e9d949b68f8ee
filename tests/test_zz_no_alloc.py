"""gqp_mala_stats_noalloc — the allocation-free variant of gqp_mala_stats (scratch in a static device array inside the
library; no counterpart in the reference).  Collected last on purpose."""
import numpy as np
import pytest


def _inputs(n_obj, B, D, seed):
    rng = np.random.default_rng(seed)
    grad = (rng.normal(size=(n_obj * B, D)) * 50).astype(np.float32)
    grad[1, 2] = np.nan
    grad[2, 3] = 1e4
    return grad, rng.normal(size=n_obj * B).astype(np.float32) * 10 + 30


def _both(lib, be, n_obj, B, D, seed, clip):
    from graspqp_b200 import _lib
    grad, energy = _inputs(n_obj, B, D, seed)
    dg, de = be.from_host(grad, np.float32), be.from_host(energy, np.float32)
    g1, g2 = be.empty((D,), np.float32), be.empty((D,), np.float32)
    z1, z2 = be.empty((n_obj * B,), np.float32), be.empty((n_obj * B,), np.float32)
    scratch = be.empty(((n_obj * B + 63) // 64 * D,), np.float32)
    p = be.ptr
    _lib.check(lib.gqp_mala_stats(p(dg), p(de), n_obj, B, D, clip, p(g1), p(z1), p(scratch), be.stream), lib)
    _lib.check(lib.gqp_mala_stats_noalloc(p(dg), p(de), n_obj, B, D, clip, p(g2), p(z2), be.stream), lib)
    H = be.to_host
    # same kernels, same reduction order: bit-identical (the parity tests' tolerance for this stage is 1e-4 / bit-exact)
    np.testing.assert_array_equal(H(g2), H(g1))
    np.testing.assert_array_equal(H(z2), H(z1))
    assert np.isfinite(H(g2)).all() if clip else True
    return dg, de, g2, z2


def _check(lib, be):
    # two calls in a row: a large batch first, then a small one that reuses (part of) the array with stale contents
    _both(lib, be, 3, 700, 25, 0, clip=1)
    _both(lib, be, 2, 65, 33, 1, clip=1)
    _both(lib, be, 1, 64, 10, 2, clip=0)
    # beyond the stated maximum: refused with a clear error, nothing launched
    from graspqp_b200 import _lib
    mx = int(lib.gqp_mala_stats_noalloc_max_floats())
    assert mx == 65536
    dg, de, g, z = _both(lib, be, 1, 64, 25, 3, clip=0)
    B_big = (mx // 25 + 1) * 64
    rc = lib.gqp_mala_stats_noalloc(be.ptr(dg), be.ptr(de), 1, B_big, 25, 0, be.ptr(g), be.ptr(z), be.stream)
    assert rc == -3 and b"static" in lib.gqp_last_error()
    with pytest.raises(_lib.GqpError):
        _lib.check(rc, lib)


def test_noalloc_stats_emulated(emu):
    from graspqp_b200 import engine
    _check(emu, engine.NumpyBackend())


@pytest.mark.gpu
def test_noalloc_stats_gpu():
    import torch
    from graspqp_b200 import _lib, engine
    assert torch.cuda.is_available()
    lib, be = _lib.lib(), engine.TorchBackend("cuda:0")
    _check(lib, be)
    torch.cuda.synchronize()
    # no allocation inside the entry point: torch's allocator statistics and the driver's free memory stay put
    grad, energy = _inputs(2, 512, 25, 5)
    dg, de = be.from_host(grad, np.float32), be.from_host(energy, np.float32)
    g, z = be.empty((25,), np.float32), be.empty((1024,), np.float32)
    torch.cuda.synchronize()
    free0, n0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_stats()["allocation.all.allocated"]
    for _ in range(3):
        _lib.check(lib.gqp_mala_stats_noalloc(be.ptr(dg), be.ptr(de), 2, 512, 25, 1, be.ptr(g), be.ptr(z), be.stream), lib)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_stats()["allocation.all.allocated"] == n0
