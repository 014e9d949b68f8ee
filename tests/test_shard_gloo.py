"""Multi-process path on the CPU: world_size 2 over gloo, the kernels run by the host emulator.

* the object partition and the final all-gather (shard.objects_of_rank / gather_results);
* a run sharded over 2 ranks with shard.GlobalStats (RMS normaliser + qpth exit-test words all-reduced per iteration)
  reproduces the single-process run over both objects;
* without GlobalStats (the default: no data-path collective) the ranks run independently and still agree with a
  single-process run of their own objects."""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from graspqp_b200 import _lib, build, engine, hand_pack, shard, synth

B, N_C, N_PTS, T, N_OBJ = 3, 12, 48, 2, 2


def _case():
    pack = hand_pack.load_pack("allegro")
    fvs, sps, verts = [], [], []
    for o in range(N_OBJ):
        v, f = synth.make_mesh("sphere" if o == 0 else "box", 200, 5 + o)
        verts.append(v)
        fvs.append(synth.face_vertices(v, f))
        sps.append(synth.sample_surface(v, f, N_PTS, seed=5 + o))
    pose, idx = synth.random_grasp_init(pack, verts, B, N_C, 9)
    pose[:, :3] *= 0.7
    rng = np.random.default_rng(1)
    noise = [(rng.random((N_OBJ * B, N_C), dtype=np.float32),
              rng.integers(0, len(pack["cand"]), (N_OBJ * B, N_C)).astype(np.int32),
              rng.random(N_OBJ * B, dtype=np.float32)) for _ in range(T)]
    return pack, fvs, np.stack(sps), pose.astype(np.float32), idx.astype(np.int32), noise


def _run(lib, objs, collectives=None):
    pack, fvs, sps, pose, idx, noise = _case()
    rows = np.concatenate([np.arange(o * B, (o + 1) * B) for o in objs])
    eng = engine.FitEngine(lib, engine.NumpyBackend(), pack, [fvs[o] for o in objs], sps[objs], B, pose[rows], idx[rows],
                           collectives=collectives)
    for us, ir, ua in noise:
        eng.step(us[rows], ir[rows], ua[rows])
    out = dict(pose=eng.pose.copy(), energy=eng.energy.copy(), idx=eng.idx.copy(), accept=eng.accept.copy(),
               stop=int(eng.qp_stop[0]))
    eng.close()
    return out


def _worker(rank, world, port, emu_path, use_stats, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lib = _lib.bind(ctypes.CDLL(emu_path), strict=False)
        objs = shard.objects_of_rank(N_OBJ, rank, world)
        stats = shard.GlobalStats() if use_stats else None
        r = _run(lib, objs, stats)
        pose, energy, idx = shard.gather_results(r["pose"], r["energy"], r["idx"], N_OBJ, B, rank, world)
        if rank == 0:
            q.put(dict(pose=pose.numpy(), energy=energy.numpy(), idx=idx.numpy(), stop=r["stop"],
                       calls=stats.calls if stats else 0))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(use_stats):
    emu_path = build.build_emu()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    mp.spawn(_worker, args=(2, _free_port(), emu_path, use_stats, q), nprocs=2, join=True)
    return q.get()


def test_partition():
    assert shard.objects_of_rank(8, 1, 4) == [1, 5]
    assert shard.objects_of_rank(3, 2, 4) == [2] and shard.objects_of_rank(3, 3, 4) == []
    assert sorted(sum((shard.objects_of_rank(11, r, 4) for r in range(4)), [])) == list(range(11))
    assert all(shard.owner_of_object(o, 4) == r for r in range(4) for o in shard.objects_of_rank(11, r, 4))
    with pytest.raises(ValueError):
        shard.objects_of_rank(4, 4, 4)


@pytest.mark.timeout(900)
def test_sharded_with_global_stats_equals_single_process(emu):
    want = _run(emu, [0, 1])
    got = _spawn(use_stats=True)
    assert got["calls"] == 2 * T + 1  # rms + qp words per iteration, qp words once at initialisation
    assert got["stop"] == want["stop"]
    np.testing.assert_array_equal(got["idx"], want["idx"])
    np.testing.assert_allclose(got["energy"], want["energy"], rtol=1e-5)
    np.testing.assert_allclose(got["pose"], want["pose"], rtol=1e-5, atol=1e-6)


@pytest.mark.timeout(900)
def test_sharded_default_runs_objects_independently(emu):
    got = _spawn(use_stats=False)
    for o in range(N_OBJ):
        want = _run(emu, [o])
        sl = slice(o * B, (o + 1) * B)
        np.testing.assert_array_equal(got["idx"][sl], want["idx"])
        np.testing.assert_allclose(got["energy"][sl], want["energy"], rtol=1e-6)
        np.testing.assert_allclose(got["pose"][sl], want["pose"], rtol=1e-6, atol=1e-7)
