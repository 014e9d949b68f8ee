#!/usr/bin/env python
"""bench.py — grasp-steps/s of the GraspQP hot path (per-iteration energy + MALA* update), BASELINE.json's metric.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload config2] [--impl reference]

A step = one iteration of scripts/fit.py:399-458 over the whole batch of this rank (N=1 workload: config 2 = Allegro,
8 synthetic 10k-triangle meshes, 1024 candidates per object, n_contact 12, 2500 surface samples).  N>1: one process per
GPU (torchrun), every rank optimises its own 8 objects (weak scaling, no data-path collective; NCCL only for the
barrier, the max-over-ranks time and the final all-gather of energies).
`--impl reference` times the CPU restatement of the reference path (oracle/, torch CPU + host-compiled TorchSDF math) on
a bounded sample of the same workload; rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "grasp-steps/sec"
UNIT = "grasp-steps/s"


def ncu_traffic_bytes(kernels):
    """dram__bytes_read + dram__bytes_write per launch of the named kernels, from the committed `ncu --set full`
    summary of this workload (profiles/r1d_ncu_full_top.txt, config 2); None when the summary is absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1d_ncu_full_top.txt")
    if not os.path.exists(path):
        return None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, cur, seen = 0.0, None, set()
    for ln in open(path):
        t = ln.split()
        if ln.startswith("=="):
            cur = next((k for k in kernels if k in ln), None)   # "== void pen_collect_kernel<1>" -> pen_collect_kernel
        elif cur in kernels and len(t) == 3 and t[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            total += float(t[1]) * unit.get(t[2], 1.0)
            seen.add(cur)
    return total if len(seen) == len(kernels) else None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="config2")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="override candidates per object")
    ap.add_argument("--objects", type=int, default=None, help="override objects per GPU")
    ap.add_argument("--no-qp-batch-stop", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="do not replay the energy evaluation from a CUDA graph")
    ap.add_argument("--no-overlap", action="store_true", help="run the penetration chain on the main stream too")
    ap.add_argument("--global-rms", action="store_true",
                    help="N>1: all-reduce the batch-wide MALA*/qpth statistics every iteration (shard.GlobalStats); the "
                         "default is the reference's practice of one independent process per object group")
    ap.add_argument("--cpu-sample-batch", type=int, default=16)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md's clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stamps, self.proc = index, [], [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])
            self.stamps.append(time.perf_counter())

    def count_since(self, t0):
        return sum(1 for t in list(self.stamps) if t >= t0)

    def stop(self, t0=0.0):
        """Summary of the rows that arrived since t0 (the start of the timed region)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in [r for r, t in zip(list(self.rows), list(self.stamps)) if t >= t0]:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(args, steps, warmup, sample_batch):
    """The CPU restatement of the reference path (oracle/) on a bounded sample: 1 object of the workload,
    `sample_batch` candidates, all host threads."""
    import torch
    from graspqp_b200 import workload
    from oracle.pipeline import FitLoopOracle
    subprocess.check_call(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], stdout=subprocess.DEVNULL)
    from oracle import sdf as osdf
    torch.set_num_threads(os.cpu_count() or 1)   # all host threads, whatever OMP_NUM_THREADS the launcher exported
    osdf.set_threads(os.cpu_count() or 1)
    w = workload.build(args.workload, batch_each=sample_batch, n_objects=1)
    B, n_c = sample_batch, w["n_contact"]
    orc = FitLoopOracle(w["pack"], w["face_verts"], [w["surface"][0]], B, w["pose"], w["idx"].astype(np.int64),
                        dtype=torch.float32)
    rng = np.random.default_rng(0)
    n_cand = len(w["pack"]["cand"])

    def one():
        orc.iterate(torch.tensor(rng.random((B, n_c), dtype=np.float32)),
                    torch.tensor(rng.integers(0, n_cand, (B, n_c))).long(), torch.tensor(rng.random(B, dtype=np.float32)))
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    cores = torch.get_num_threads()
    return dict(value=B * steps / dt, unit=UNIT, cores=cores, kind="port", seconds=dt,
                sample=f"{args.workload}: 1 object ({w['kinds'][0]}, {w['faces_per_mesh']} tris) x {B} candidates x "
                       f"{steps} iterations, 2500 surface samples, torch-CPU fp32 + OpenMP TorchSDF math "
                       f"({os.cpu_count()} host cpus)"), dt / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to its workers; this arm is the reference's CPU path on ALL host threads
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # bounded: the oracle does ~0.15 s per iteration of 16 candidates on 16 cores
    steps, warmup = min(steps, 100), min(warmup, 3)
    cb, sec_per_step = cpu_reference_run(args, steps, warmup, args.cpu_sample_batch)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * sec_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "sample": cb["sample"]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_b200(args):
    import torch
    import torch.distributed as dist
    from graspqp_b200 import _lib, engine, shard, workload
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()
    be = engine.TorchBackend(dev)
    w = workload.build(args.workload, rank=rank, batch_each=args.batch, n_objects=args.objects)
    B, n_c, O = w["batch_each"], w["n_contact"], len(w["face_verts"])
    Bt = B * O
    t_setup = time.perf_counter()
    eng = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], B, w["pose"], w["idx"],
                           qp_batch_stop=not args.no_qp_batch_stop, seed=1 + rank,
                           collectives=shard.GlobalStats() if (world > 1 and args.global_rms) else None,
                           overlap=not args.no_overlap)
    graphed = (not args.no_graph) and eng.capture()
    torch.cuda.synchronize()
    t_setup = time.perf_counter() - t_setup
    K, W = max(1, args.steps), max(3, args.warmup)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (Philox noise drawn on the device) ----
    sampler = ClockSampler(local)
    sampler.start()  # started before the warm-up: nvidia-smi needs a few hundred ms before its first row
    for _ in range(W):
        eng.step()
    barrier()
    after_warmup = eng.snapshot()  # the end-to-end loop below times the same K iterations of the fit
    t_timed = time.perf_counter()
    n0 = lib.gqp_launch_count()
    evs = []
    for _ in range(K):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.step()
        e1.record()
        evs.append((e0, e1))
    barrier()
    # kernels of this library executed in the timed region: counted launch calls + the kernels inside each graph replay
    launches = int(lib.gqp_launch_count() - n0) + (K * eng.graph_launches if graphed else 0)
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    # K iterations can be shorter than nvidia-smi's 100 ms period: keep the GPU under the SAME load (untimed iterations,
    # the state is restored before the end-to-end loop) until two rows have arrived since the timed region began
    t_load = time.perf_counter()
    while sampler.proc is not None and sampler.count_since(t_timed) < 2 and time.perf_counter() - t_load < 3.0:
        flush.zero_()
        eng.step()
        torch.cuda.synchronize()
    clocks = sampler.stop(t_timed)
    # ---- end-to-end: host noise in pinned memory -> device, step, energy + accept back to the host, every step ----
    n_cand = len(w["pack"]["cand"])
    rng = np.random.default_rng(rank)
    pin = lambda a: torch.from_numpy(a).pin_memory()
    h_us = [pin(rng.random((Bt, n_c), dtype=np.float32)) for _ in range(K)]
    h_ir = [pin(rng.integers(0, n_cand, (Bt, n_c)).astype(np.int32)) for _ in range(K)]
    h_ua = [pin(rng.random(Bt, dtype=np.float32)) for _ in range(K)]
    # double-buffered on both sides: the noise of iteration k+1 is uploaded and iteration k+1 enqueued before the host
    # waits for iteration k, whose energies / accept flags it then reads (one iteration behind, as a logging caller
    # does) — every iteration's inputs still come from pinned host memory and every iteration's result is read back
    d_noise = [(be.empty((Bt, n_c), np.float32), be.empty((Bt, n_c), np.int32), be.empty((Bt,), np.float32))
               for _ in range(2)]
    h_out = [(torch.empty(Bt, dtype=torch.float32).pin_memory(), torch.empty(Bt, dtype=torch.uint8).pin_memory())
             for _ in range(2)]
    main, cs = torch.cuda.current_stream(), torch.cuda.Stream(device=dev)  # cs: copy stream (both directions)
    up_done, step_done, out_done = ([torch.cuda.Event() for _ in range(2)] for _ in range(3))

    def upload(k):  # on the copy stream, once the last iteration that read this buffer pair (k-2) has finished
        a, b, c = d_noise[k % 2]
        with torch.cuda.stream(cs):
            if k >= 2:
                cs.wait_event(step_done[k % 2])
            a.copy_(h_us[k], non_blocking=True); b.copy_(h_ir[k], non_blocking=True); c.copy_(h_ua[k], non_blocking=True)
            up_done[k % 2].record(cs)

    host_sum = 0.0
    eng.restore(after_warmup)
    barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    upload(0)
    for k in range(K):
        main.wait_event(up_done[k % 2])
        if k >= 1:
            main.wait_event(out_done[(k - 1) % 2])  # the previous result has left eng.energy / eng.accept
        eng.step(*d_noise[k % 2])
        step_done[k % 2].record(main)
        with torch.cuda.stream(cs):
            cs.wait_event(step_done[k % 2])
            h_out[k % 2][0].copy_(eng.energy, non_blocking=True); h_out[k % 2][1].copy_(eng.accept, non_blocking=True)
            out_done[k % 2].record(cs)
        if k + 1 < K:
            upload(k + 1)
        if k > 0:
            out_done[(k - 1) % 2].synchronize()
            host_sum += float(h_out[(k - 1) % 2][0][0])  # the caller touches the result of every iteration
    out_done[(K - 1) % 2].synchronize()
    host_sum += float(h_out[(K - 1) % 2][0][0])
    s1.record()
    barrier()
    ms_e2e = s0.elapsed_time(s1)
    h2d = Bt * n_c * 8 + Bt * 4
    d2h = Bt * 5
    # ---- every stage alone on the current state (CUDA events on the launching stream, L2 flushed before each) ----
    stage_ms = {}
    for name, fn in eng.stage_calls().items():
        for _ in range(2):
            _lib.check(fn(), lib)
        ts = []
        for _ in range(max(5, min(K, 10))):
            flush.zero_()
            k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            k0.record()
            _lib.check(fn(), lib)
            k1.record()
            torch.cuda.synchronize()
            ts.append(k0.elapsed_time(k1))
        stage_ms[name] = float(np.median(ts))
    ms_pen = stage_ms["hand_distance"]
    # FP32 FMA peak, measured live
    sink = be.empty((4,), np.float32)
    import ctypes
    nf = ctypes.c_uint64(0)
    lib.gqp_fma_peak_probe(2000, be.ptr(sink), ctypes.byref(nf), be.stream)
    torch.cuda.synchronize()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    lib.gqp_fma_peak_probe(20000, be.ptr(sink), ctypes.byref(nf), be.stream)
    f1.record()
    torch.cuda.synchronize()
    fp32_peak_tflops = 2.0 * nf.value / (f0.elapsed_time(f1) * 1e-3) / 1e12
    # ---- max over ranks, gather ----
    t = torch.tensor([ms_total, ms_e2e, ms_pen], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # final grasps of all ranks in global object order (the only data exchange of the path by default)
        shard.gather_results(eng.pose, eng.energy, eng.idx, O * world, B, rank, world)
    ms_total, ms_e2e, ms_pen = [float(v) for v in t.tolist()]
    final_energy = float(eng.energy.float().mean().item())
    if rank == 0:
        f_hand = int(w["pack"]["face_offset"][-1])
        pen_flops = 80.0 * eng.n_pts * f_hand * Bt  # brute-force-equivalent point-face tests x 80 flop
        achieved = pen_flops / (ms_pen * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        alg_bytes = Bt * (eng.L * 48 + 4 + 4 * ((eng.n_pts + 31) // 32)) + O * eng.n_pts * 12
        line = {
            "metric": METRIC, "value": Bt * world * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['hand']}, {O} meshes/GPU ({','.join(w['kinds'])}; "
                                   f"{w['faces_per_mesh']} tris), batch {B}/object, n_contact {n_c}, 2500 surface samples",
                       "candidates_per_gpu": Bt, "l2": "flushed between timed steps (256 MiB write)",
                       "qp_batch_stop": eng.qp_batch_stop, "streams": 2 if eng.overlap else 1, "cuda_graph": bool(graphed),
                       "collectives": ("global rms + qp exit words all-reduced per step" if eng.collectives else
                                       "none on the data path; final all-gather of grasps") if world > 1 else "n/a",
                       "noise": "philox on device (value) / injected from host (e2e)",
                       "e2e_pipeline": "double-buffered, copies on a second stream: upload k+1 and enqueue k+1 before reading back k",
                       "setup_s": round(t_setup, 2), "final_mean_energy": final_energy},
            "e2e": {"value": Bt * world * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "clocks": clocks,
            "stage_ms": {k: round(v, 4) for k, v in stage_ms.items()},
            "roofline": {"bound": "fp32", "kernel": "H4 hand penetration = pen_collect_kernel + pen_search_kernel "
                                                    "(98.5 % of a grasp-step's brute-force-equivalent work)",
                         "achieved": achieved,
                         "peak": fp32_peak_tflops, "unit": "TFLOP/s", "frac": achieved / fp32_peak_tflops,
                         "traffic": ncu_traffic_bytes(("pen_broad_kernel", "pen_collect_kernel", "pen_search_kernel")),
                         "ms_per_launch": ms_pen,
                         "note": "achieved = brute-force-equivalent flops (80 x n_pts x F_hand x candidates) / launch "
                                 "time; exact culling skips most pairs, so frac may exceed 1 (SURVEY 8d). peak = FP32 "
                                 "FMA probe measured live in this run",
                         "hbm": {"achieved": alg_bytes / (ms_pen * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": alg_bytes / (ms_pen * 1e-3) / 1e9 / hbm_peak,
                                 "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}},
        }
        if world == 1 and not args.skip_cpu_baseline:
            cb, _ = cpu_reference_run(args, 100, 1, args.cpu_sample_batch)   # ~15 s of host work
            line["cpu_baseline"] = cb
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
