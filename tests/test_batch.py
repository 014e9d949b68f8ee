"""(f4) batch driver (graspqp_b200/batch.py): the queue / skip-if-exists / batching rules of scripts/fit_all.sh, sharding of
the batches over ranks (world size 2, gloo), entropy statistics, and one tiny batch end to end through the emulated kernels."""
import os

import numpy as np
import pytest
import torch

from graspqp_b200 import batch, engine, hand_pack, synth


def _write_obj(path, v, f):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as fh:
        for p in v:
            fh.write("v %.9g %.9g %.9g\n" % tuple(p))
        for t in f:
            fh.write("f %d %d %d\n" % tuple(int(i) + 1 for i in t))


def _dataset(root, names, done=()):
    for k, n in enumerate(names):
        v, f = synth.make_mesh(["sphere", "box", "cylinder"][k % 3], 300, k)
        _write_obj(os.path.join(root, n, "coacd", "decomposed.obj"), v, f)
    for n in done:  # fit_all.sh:160: the file whose presence skips the asset
        p = batch.result_file(root, n, "allegro", 12, "graspqp", "default", 7000)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        open(p, "wb").close()
    open(os.path.join(root, "not_a_directory.txt"), "w").close()


def test_queue_skip_and_batches(tmp_path):
    root = str(tmp_path)
    names = [f"asset_{k:02d}" for k in range(7)]
    _dataset(root, names, done=("asset_02", "asset_05"))
    queue, skipped = batch.scan_assets(root, "allegro", 12, "graspqp")
    assert skipped == ["asset_02", "asset_05"] and queue == [n for n in names if n not in skipped]
    # a result of another hand / contact count / step does not skip (fit_all.sh:160 path rule)
    assert batch.scan_assets(root, "shadow_hand", 12, "graspqp")[1] == []
    assert batch.scan_assets(root, "allegro", 8, "graspqp")[1] == []
    assert batch.scan_assets(root, "allegro", 12, "graspqp", step=100)[1] == []
    b = batch.make_batches(queue, 2)
    assert b == [["asset_00", "asset_01"], ["asset_03", "asset_04"], ["asset_06"]]  # remainder batch: fit_all.sh:217-243
    assert batch.make_batches(queue, 64) == [queue] and batch.make_batches([], 4) == []
    with pytest.raises(ValueError):
        batch.make_batches(queue, 0)
    with pytest.raises(FileNotFoundError):
        batch.scan_assets(os.path.join(root, "missing"), "allegro", 12, "graspqp")
    # every batch on exactly one rank
    for world in (1, 2, 3, 8):
        got = sorted(k for r in range(world) for k, _ in batch.batches_of_rank(b, r, world))
        assert got == list(range(len(b)))
    assert batch.main(["-d", root, "-f", "", "-a", "2", "--dry-run"]) == 0


def test_mesh_path_rule_and_obj_reader(tmp_path):
    root = str(tmp_path)
    v, f = synth.make_mesh("sphere", 300, 0)
    _write_obj(os.path.join(root, "a", "coacd", "decomposed.obj"), v, f)
    assert batch.find_mesh(root, "a").endswith("decomposed.obj")
    _write_obj(os.path.join(root, "a", "coacd", "remeshed.obj"), v, f)
    assert batch.find_mesh(root, "a").endswith("remeshed.obj")  # object_model.py:97-99: remeshed first
    _write_obj(os.path.join(root, "b", "model.obj"), v, f)
    assert batch.find_mesh(root, "b").endswith("model.obj")      # :100-113: any *.obj of the directory
    with pytest.raises(FileNotFoundError):
        batch.find_mesh(root, "c")
    v2, f2 = batch.load_obj(batch.find_mesh(root, "a"))
    np.testing.assert_allclose(v2, v, rtol=1e-8)
    np.testing.assert_array_equal(f2, f)
    with open(os.path.join(root, "quad.obj"), "w") as fh:  # polygons are fan-triangulated; too few vertices raise
        fh.write("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3 4\n")
    with pytest.raises(ValueError):
        batch.load_obj(os.path.join(root, "quad.obj"))


def _rank_main(rank, world, root, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    queue, _ = batch.scan_assets(root, "allegro", 12, "graspqp")
    mine = [k for k, _ in batch.batches_of_rank(batch.make_batches(queue, 2), rank, world)]
    t = torch.zeros(8, dtype=torch.int64)
    t[mine] = 1
    dist.all_reduce(t)  # test plumbing only: the driver itself has no collective
    if rank == 0:
        torch.save(t, out)
    dist.destroy_process_group()


def test_every_batch_taken_once_world2(tmp_path):
    import torch.multiprocessing as mp
    root = str(tmp_path / "data")
    _dataset(root, [f"asset_{k:02d}" for k in range(7)], done=("asset_03",))
    out = str(tmp_path / "taken.pt")
    mp.spawn(_rank_main, args=(2, root, 29571, out), nprocs=2, join=True)
    assert torch.load(out).tolist() == [1, 1, 1, 0, 0, 0, 0, 0]  # 6 queued assets -> 3 batches, each on one rank


def test_entropy_statistics(emu):
    """gqp_entropy_stats vs the restatement of HandModel.joint_entropy / pose_entropy with the real torch.histc"""
    from oracle import entropy as oe
    pack = hand_pack.load_pack("allegro")
    v, f = synth.make_mesh("sphere", 200, 3)
    B = 300
    pose, idx = synth.random_grasp_init(pack, [v], B, 12, 3)
    pose[:40, 9] = pack["joints_upper"][0]   # on the upper limit: last bin
    pose[40:60, 10] = 9.0                    # outside the limits: dropped
    pose[60:70, 0] = 0.5                     # translation outside +-0.1: dropped
    # the entry point itself (FitEngine.entropy wraps exactly this call)
    pose = np.ascontiguousarray(pose, np.float32)
    lo, hi = np.ascontiguousarray(pack["joints_lower"], np.float32), np.ascontiguousarray(pack["joints_upper"], np.float32)
    dof = len(lo)
    scratch, out = np.zeros(dof + 6, np.float32), np.zeros(3, np.float32)
    ptr = lambda a: a.ctypes.data
    assert emu.gqp_entropy_stats(ptr(pose), B, dof, ptr(lo), ptr(hi), ptr(scratch), ptr(out), None) == 0
    e = {"joints_entropy": float(out[0]), "translation_entropy": float(out[1]), "rotation_entropy": float(out[2]),
         "total": float(0.5 * out[0] + 0.5 * (out[1] + out[2]))}
    hp = torch.tensor(pose, dtype=torch.float32)
    je = float(oe.joint_entropy(hp, torch.tensor(pack["joints_lower"]), torch.tensor(pack["joints_upper"])))
    te, re_ = [float(x) for x in oe.pose_entropy(hp)]
    assert abs(e["joints_entropy"] - je) < 2e-3 and abs(e["translation_entropy"] - te) < 2e-3  # a value on a bin edge may
    assert abs(e["rotation_entropy"] - re_) < 2e-3                                            # fall either side: 1/300 rows
    assert abs(e["total"] - (0.5 * je + 0.5 * (te + re_))) < 3e-3


def test_fit_batch_end_to_end_emulated(emu, tmp_path):
    """two small objects, 3 candidates each, 3 iterations: meshes read from disk, device initialisation, fit, export to the
    reference's path, then the driver skips them"""
    root = str(tmp_path)
    names = ["obj_a", "obj_b"]
    _dataset(root, names)
    r = batch.fit_batch(root, names, hand="allegro", n_contact=4, batch_size=3, n_iter=3, n_surface=48, reset_epochs=0,
                        log_entropy=True, cdll=emu, backend=engine.NumpyBackend())
    assert len(r["paths"]) == 2 and np.isfinite(r["mean_energy"]) and r["entropy"]["total"] >= 0
    for n, p in zip(names, r["paths"]):
        assert p == batch.result_file(root, n, "allegro", 4, "graspqp", "default", 3)
        d = torch.load(p, weights_only=False)
        assert d["values"].shape == (3,) and d["parameters"]["root_pose"].shape == (3, 7) and d["contact_idx"].shape == (3, 4)
    queue, skipped = batch.scan_assets(root, "allegro", 4, "graspqp", step=3)
    assert queue == [] and skipped == names
