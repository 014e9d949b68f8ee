"""BASELINE G — the reference's GPU path on the same B200: the loop body of scripts/fit.py:399-458
run with the reference's OWN code wherever it can run in this image, on CUDA tensors:

    MalaStar.try_step / accept_step / zero_grad     baseline/_ref/graspqp/core/optimizer.py      (unmodified)
    calculate_energy                                baseline/_ref/graspqp/core/energy.py         (unmodified)
    GraspSpanMetricFactory / span / SQPLsqSolver    baseline/_ref/graspqp/metrics/...            (unmodified)
    torchsdf.compute_sdf (the reference's one native kernel, compiled unmodified for sm_100)     baseline/_ref/torchsdf

What is NOT the reference (the packages are absent from this image and cannot be installed: no network):
    qpth.qp.QPFunction      -> oracle.qp.QPFunction on CUDA tensors (restatement of qpth 0.0.18's batched PDIPM: the same
                               algorithm and the same batched LU of the 2n x 2n Schur complement, torch.linalg on cuSOLVER)
    HandModel / ObjectModel -> oracle.hand.HandOracle / oracle.pipeline.ObjectOracle (trimesh, pytorch3d, lxml absent):
                               op-by-op restatements of hand_model.py:787-1040 / object_model.py:186-255 in torch, calling
                               the real torchsdf.compute_sdf once per link / object exactly as the reference does.
So the number is "reference algorithm and launch structure, reference SDF kernel, stock PyTorch autograd" — the
denominator north_star's ">= 50x" target names — with those two substitutions stated.

    python baseline/gpu_reference.py [--workload config2] [--batch B] [--objects O] [--steps K] [--warmup W]
prints one JSON object.  Run by bench.py (`gpu_baseline` key) in a subprocess; never imported by graspqp_b200/.
"""
import argparse
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))
REFDIR = os.path.join(ROOT, "baseline", "_ref")


def available():
    return os.path.exists(os.path.join(REFDIR, "torchsdf", "_C.so")) and os.path.isdir(os.path.join(REFDIR, "graspqp"))


def run(workload_name="config2", batch_each=None, n_objects=None, steps=5, warmup=2, device="cuda:0", seed=1):
    import numpy as np
    import torch
    import ref_loader
    from graspqp_b200 import workload
    from oracle.hand import HandOracle
    from oracle.pipeline import ObjectOracle, WEIGHTS
    sys.path.insert(0, REFDIR)
    import torchsdf  # the reference's extension, built by baseline/build_ref_gpu.sh
    ref = ref_loader.load(os.path.join(REFDIR, "graspqp"), os.path.join(REFDIR, "pytorch_kinematics"))
    dev = torch.device(device)
    torch.manual_seed(seed)  # fit.py:159-160
    w = workload.build(workload_name, batch_each=batch_each, n_objects=n_objects)
    B, O, n_c = w["batch_each"], len(w["face_verts"]), w["n_contact"]
    dtype = torch.float32
    hand = HandOracle(w["pack"], dtype, device=dev, sdf=torchsdf.compute_sdf)
    obj = ObjectOracle(w["face_verts"], list(w["surface"]), B, dtype, device=dev, sdf=torchsdf.compute_sdf)
    obj.device = dev
    obj.object_scale_tensor = torch.ones(O, B, dtype=dtype, device=dev)
    hand.set_parameters(torch.as_tensor(w["pose"], dtype=dtype, device=dev).clone().requires_grad_(True),
                        torch.as_tensor(w["idx"], device=dev).long())
    opt = ref["optimizer"].MalaStar(hand, device=dev, batch_size=B, switch_possibility=0.4, starting_temperature=18.0,
                                    temperature_decay=0.95, annealing_period=30, step_size=0.005, stepsize_period=50,
                                    mu=0.98)  # scripts/fit.py argparse defaults
    F = ref["registry"].GraspSpanMetricFactory
    energy_fnc = F.create(F.MetricType.GRASPQP, solver_kwargs={"friction": 0.2, "max_limit": 20.0, "n_cone_vecs": 4})
    names = [k for k, v in WEIGHTS.items() if v > 0]

    def energies():
        losses = ref["energy"].calculate_energy(hand, obj, energy_fnc=energy_fnc, energy_names=names,
                                                method="gendexgrasp", svd_gain=0.1)
        e = 0
        for k, v in losses.items():
            e = e + WEIGHTS[k] * v
        return losses, e

    losses, energy = energies()  # fit.py:381-396
    energy.sum().backward()
    opt.zero_grad()
    energy = energy.detach()
    losses = {k: v.detach().clone() for k, v in losses.items()}

    def iterate():  # fit.py:399-458 (no reset: same as the engine's timed iterations)
        nonlocal energy
        opt.try_step()
        Eb = energy.view(-1, B)
        z = ((Eb - Eb.mean(-1).unsqueeze(-1)) / Eb.std(-1).unsqueeze(-1)).view(-1)
        opt.zero_grad()
        new_losses, new_energy = energies()
        new_energy.sum().backward()
        with torch.no_grad():
            accept, _ = opt.accept_step(energy, new_energy, None, z, 1.0)
            energy[accept] = new_energy[accept]
            for k, v in new_losses.items():
                losses[k][accept] = v[accept]

    for _ in range(warmup):
        iterate()
    torch.cuda.synchronize(dev)
    evs = []
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        iterate()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    return dict(value=O * B / (ms * 1e-3), unit="grasp-steps/s", ms_per_step=ms, steps=steps, warmup=warmup,
                candidates=O * B, workload=f"{workload_name}: {w['hand']}, {O} meshes x {w['faces_per_mesh']} tris, batch {B}/object, "
                                           f"n_contact {n_c}, 2500 surface samples, fp32",
                kind="reference GPU path: reference MalaStar / calculate_energy / span / qp_solver + reference TorchSDF "
                     "kernel (sm_100 build) + torch autograd; qpth -> oracle.qp on CUDA (cuSOLVER batched LU), "
                     "HandModel/ObjectModel -> torch restatements calling torchsdf once per link/object",
                final_mean_energy=float(energy.mean().item()))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config2")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--objects", type=int, default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    if not available():
        print(json.dumps({"unavailable": "baseline/_ref not built (run baseline/build_ref_gpu.sh where /root/reference exists)"}))
    else:
        print(json.dumps(run(a.workload, a.batch, a.objects, a.steps, a.warmup)))
