"""(f4) Batch driver — what scripts/fit_all.sh + scripts/fit.py do around the loop body, for ONE launcher feeding N GPUs.

The reference scans a dataset folder, skips assets whose result file exists (fit_all.sh:154-168), cuts the rest into batches of
`num_assets` objects (:170-214), writes one run script per batch and leaves it to the cluster scheduler to run each as its own
scripts/fit.py process on one GPU (:251-255).  Here the same queue is consumed by the ranks of one torchrun job: batch k goes to
rank k mod world (no data-path collective: batches are independent); every rank then runs

    load meshes (object_model.py:97-136 path rule)  ->  surface samples  ->  FitEngine + device initialisation
    ->  fit(n_iter) with outlier resets (fit.py:399-458)  ->  entropy statistics (fit.py:462-471)  ->  export_poses (fit.py:224-300)

    torchrun --nproc-per-node 8 -m graspqp_b200.batch --data-root /data --dataset release/debug --hand allegro

Host logic only (paths, queue, sharding, OBJ reading); every number comes from libgraspqp_b200.so through FitEngine.
"""
import argparse
import glob
import os

import numpy as np

N_ITER_DEFAULT = 7000  # scripts/fit.py: --n_iter; the step the skip test looks for (fit_all.sh:160)


def result_dir(data_root_path, object_code, hand, n_contact, energy_name, grasp_type="default"):
    """scripts/fit.py:203-221 (get_result_path)"""
    return os.path.join(data_root_path, object_code, "grasp_predictions", hand, f"{n_contact}_contacts", energy_name,
                        "default" if grasp_type in (None, "all") else grasp_type)


def result_file(data_root_path, object_code, hand, n_contact, energy_name, grasp_type="default", step=N_ITER_DEFAULT):
    """fit_all.sh:160 (CHECK_FILE_PATH): <asset>/grasp_predictions/<hand>/<n>_contacts/<energy>/<type>/<asset>_step_<n>.dexgrasp.pt"""
    return os.path.join(result_dir(data_root_path, object_code, hand, n_contact, energy_name, grasp_type),
                        f"{object_code}_step_{step}.dexgrasp.pt")


def scan_assets(folder, hand, n_contact, energy_name, grasp_type="default", step=N_ITER_DEFAULT):
    """fit_all.sh:154-168: the asset directories of `folder` in sorted order (bash globbing sorts) minus those whose result
    file exists.  -> (queue, skipped)"""
    if not os.path.isdir(folder):
        raise FileNotFoundError(f"Folder '{folder}' does not exist!")  # fit_all.sh:121-124
    queue, skipped = [], []
    for name in sorted(os.listdir(folder)):
        if not os.path.isdir(os.path.join(folder, name)):
            continue
        (skipped if os.path.isfile(result_file(folder, name, hand, n_contact, energy_name, grasp_type, step))
         else queue).append(name)
    return queue, skipped


def make_batches(queue, num_assets):
    """fit_all.sh:170-214 + 217-243: consecutive groups of num_assets assets, the remainder as a last, smaller batch"""
    if num_assets <= 0:
        raise ValueError("Number of assets must be a positive integer!")  # fit_all.sh:136-139
    return [queue[i:i + num_assets] for i in range(0, len(queue), num_assets)]


def batches_of_rank(batches, rank, world):
    """batch k runs on rank k mod world; -> list of (k, batch)"""
    return [(k, b) for k, b in enumerate(batches) if k % world == rank]


# ---- object data (object_model.py:97-136) ----------------------------------------------------------------------------
def find_mesh(data_root_path, object_code, extension=".obj"):
    for name in ("remeshed.obj", "decomposed.obj"):
        p = os.path.join(data_root_path, object_code, "coacd", name)
        if os.path.exists(p):
            return p
    meshes = sorted(glob.glob(os.path.join(data_root_path, object_code, f"*{extension}")))
    remeshed = [m for m in meshes if "remeshed.obj" in m]
    if len(remeshed) == 1:
        return remeshed[0]
    if not meshes:
        raise FileNotFoundError(f"Object {object_code} not found under {data_root_path}")
    return meshes[0]


def load_obj(path):
    """Wavefront OBJ -> (vertices (V,3) float64, triangles (F,3) int64); polygons are fan-triangulated.  (The reference
    reads with trimesh.load(force="mesh", process=True), which also merges duplicate vertices: irrelevant for the face-vertex
    table the kernels take.)"""
    verts, faces = [], []
    with open(path, "r", errors="ignore") as fh:
        for line in fh:
            if line.startswith("v "):
                verts.append([float(x) for x in line.split()[1:4]])
            elif line.startswith("f "):
                idx = [int(tok.split("/")[0]) for tok in line.split()[1:]]
                idx = [i - 1 if i > 0 else len(verts) + i for i in idx]
                for k in range(1, len(idx) - 1):
                    faces.append([idx[0], idx[k], idx[k + 1]])
    v, f = np.asarray(verts, np.float64).reshape(-1, 3), np.asarray(faces, np.int64).reshape(-1, 3)
    if len(v) < 100:
        raise ValueError(f"Object {path} has too few vertices, please check the data.")  # object_model.py:117-118
    return v, f


def load_objects(data_root_path, object_codes, n_surface=2500, seed=1):
    """-> (face_verts_list, surface_points (O, n_surface, 3), hull_tris_list) for FitEngine / enable_device_init"""
    from . import initializations, synth
    fvs, sps, hulls = [], [], []
    for k, code in enumerate(object_codes):
        v, f = load_obj(find_mesh(data_root_path, code))
        fvs.append(synth.face_vertices(v, f))
        sps.append(synth.sample_surface(v, f, n_surface, seed=seed + k))
        hulls.append(initializations.convex_hull(v))
    return fvs, np.stack(sps), hulls


# ---- one batch on one device -----------------------------------------------------------------------------------------
def fit_batch(data_root_path, object_codes, hand="allegro", n_contact=12, batch_size=32, n_iter=N_ITER_DEFAULT,
              energy_type="graspqp", optimizer="mala_star", grasp_type="default", device="cuda:0", seed=1, reset_epochs=600,
              z_score_threshold=1.0, n_surface=2500, log_entropy=False, cdll=None, backend=None, **engine_kw):
    """scripts/fit.py for the objects of one batch.  -> dict(paths, mean_energy, entropy)"""
    from . import _lib, engine, export, hand_pack, initializations
    lib = cdll if cdll is not None else _lib.lib()
    be = backend if backend is not None else engine.TorchBackend(device)
    pack = hand_pack.load_pack(hand)
    fvs, sps, hulls = load_objects(data_root_path, object_codes, n_surface, seed)
    # first poses: initialize_convex_hull on the device (core/initializations.py:15-193), then the engine takes them over
    D = 9 + len(pack["joints_lower"])
    Bt = len(object_codes) * batch_size
    init = initializations.DeviceInit(lib, be, pack, hulls, batch_size, n_contact)
    hand_handle = hand_pack.create_hand(lib, pack)
    pose0, idx0 = be.empty((Bt, D), np.float32), be.empty((Bt, n_contact), np.int32)
    init.run(hand_handle, pose0, idx0, None, None, None, seed=seed)
    pose_h, idx_h = be.to_host(pose0), be.to_host(idx0)
    init.close()
    lib.gqp_hand_destroy(hand_handle)
    eng = engine.FitEngine(lib, be, pack, fvs, sps, batch_size, pose_h, idx_h, seed=seed, energy_type=energy_type,
                           optimizer=optimizer, **engine_kw)
    eng.enable_device_init(hulls)
    eng.fit(n_iter, reset_epochs=reset_epochs, z_score_threshold=z_score_threshold)
    eng.assert_ok()
    out = {"mean_energy": float(be.to_host(eng.energy).mean()), "entropy": eng.entropy() if log_entropy else None}
    out["paths"] = export.Exporter(eng).export_poses(data_root_path, list(object_codes), hand, energy_name=energy_type,
                                                     suffix=f"_step_{n_iter}", grasp_type=grasp_type)
    eng.close()
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("-d", "--data-path", default="/data")                    # fit_all.sh:4-11 defaults
    ap.add_argument("-f", "--folder-suffix", default="release/debug")
    ap.add_argument("--hand", default="allegro")
    ap.add_argument("-g", "--grasp-type", default="default")
    ap.add_argument("-e", "--energy-method", default="graspqp")
    ap.add_argument("-c", "--n-contacts", type=int, default=12)
    ap.add_argument("-n", "--n-grasps-per-obj", type=int, default=32)
    ap.add_argument("-a", "--num-assets", type=int, default=64)
    ap.add_argument("--n-iter", type=int, default=N_ITER_DEFAULT)
    ap.add_argument("--optimizer", default="mala_star", choices=["mala_star", "dexgraspnet"])
    ap.add_argument("--log-entropy", action="store_true")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--dry-run", action="store_true", help="print the queue and the batches of this rank, fit nothing")
    a = ap.parse_args(argv)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    folder = os.path.join(a.data_path, a.folder_suffix)
    queue, skipped = scan_assets(folder, a.hand, a.n_contacts, a.energy_method, a.grasp_type, a.n_iter)
    batches = make_batches(queue, a.num_assets)
    mine = batches_of_rank(batches, rank, world)
    print(f"[rank {rank}/{world}] {len(queue)} assets queued, {len(skipped)} skipped (already processed), "
          f"{len(batches)} batches, {len(mine)} on this rank")
    if a.dry_run:
        for k, b in mine:
            print(f"[rank {rank}] batch {k}: {' '.join(b)}")
        return 0
    device = f"cuda:{int(os.environ.get('LOCAL_RANK', 0))}"
    for k, b in mine:
        r = fit_batch(folder, b, hand=a.hand, n_contact=a.n_contacts, batch_size=a.n_grasps_per_obj, n_iter=a.n_iter,
                      energy_type=a.energy_method, optimizer=a.optimizer, grasp_type=a.grasp_type, device=device,
                      seed=a.seed + k, log_entropy=a.log_entropy)
        print(f"[rank {rank}] batch {k}: mean energy {r['mean_energy']:.4f}, {len(r['paths'])} files"
              + (f", entropy {r['entropy']}" if r["entropy"] else ""))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
