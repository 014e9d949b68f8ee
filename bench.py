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


COUNTER_METRICS = [  # collected by an ncu subprocess of THIS run (never from a file): instruction counts and DRAM bytes
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
    "smsp__thread_inst_executed.sum", "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__time_duration.sum"]
# stage -> the kernels one call of it launches (names as ncu prints them, prefix match)
STAGE_KERNELS = {
    "object_distance": ("object_distance_kernel",),
    "fc_solve": ("fc_dis_kernel<1,", "fc_dis_kernel<(int)1,"),  # <MODE = FC_SOLVE, candidates per warp>
    "hand_distance": ("pen_bounds_kernel", "pen_broad_kernel", "pen_collect_kernel", "pen_search_kernel",
                      "pen_reduce_fixed_kernel"),
    "span_energy": ("span_alt_kernel",),  # --energy-type dexgrasp | tdg: replaces fc_solve + fc_finish
}


def stages_of(energy_type):
    return [st for st in STAGE_KERNELS if (st == "span_energy") == (energy_type != "graspqp") or
            st in ("object_distance", "hand_distance")]


def collect_counters(args, at_iteration):
    """Runs `bench.py --count-mode` under ncu (--metrics above, cudaProfilerStart/Stop around one call of each stage at
    iteration `at_iteration` of the same seeded fit) and sums the metrics per stage.  Counts and bytes only — every TIME
    in the bench line comes from CUDA events of the un-profiled run.  -> (dict stage -> metric sums, note)"""
    import csv
    import shutil
    import tempfile
    ncu = shutil.which("ncu") or "/usr/local/cuda/bin/ncu"
    if not os.path.exists(ncu):
        return None, "ncu not found"
    out = tempfile.NamedTemporaryFile(suffix=".csv", delete=False).name
    cmd = [ncu, "--metrics", ",".join(COUNTER_METRICS), "--clock-control", "none", "--profile-from-start", "off", "--csv",
           "--log-file", out, sys.executable, os.path.abspath(__file__), "--count-mode", "--workload", args.workload,
           "--count-at", str(at_iteration), "--energy-type", args.energy_type, "--optimizer", args.optimizer]
    if args.batch:
        cmd += ["--batch", str(args.batch)]
    if args.objects:
        cmd += ["--objects", str(args.objects)]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    except (subprocess.TimeoutExpired, OSError) as e:
        return None, f"ncu subprocess failed: {type(e).__name__}"
    if r.returncode != 0 or not os.path.exists(out):
        return None, f"ncu exited {r.returncode}: {(r.stdout + r.stderr)[-200:]!r}"
    rows = list(csv.reader(open(out, errors="replace")))
    hdr = next((k for k, row in enumerate(rows) if row and row[0] == "ID"), None)
    if hdr is None:
        return None, "ncu wrote no launch table"
    H = rows[hdr]
    ki, mi, vi = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value")
    sums = {st: {} for st in stages_of(args.energy_type)}
    for row in rows[hdr + 1:]:
        if len(row) <= vi:
            continue
        name = row[ki].replace("void ", "")
        for st in sums:
            ks = STAGE_KERNELS[st]
            if any(name.startswith(k) for k in ks):
                try:
                    sums[st][row[mi]] = sums[st].get(row[mi], 0.0) + float(row[vi].replace(",", ""))
                except ValueError:
                    pass
    os.unlink(out)
    if not all(sums.values()):
        return None, "ncu table lacks the stage kernels"
    return sums, f"ncu --metrics subprocess of this run, one call of each stage at iteration {at_iteration}"


def run_count_mode(args):
    """Child of collect_counters: the same seeded fit up to --count-at, then ONE call of every stage between
    cudaProfilerStart/Stop.  Prints nothing the parent needs; ncu's csv is the output."""
    import torch
    from graspqp_b200 import _lib, engine, workload
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib, be = _lib.lib(), engine.TorchBackend(dev)
    w = workload.build(args.workload, batch_each=args.batch, n_objects=args.objects)
    eng = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], w["batch_each"], w["pose"], w["idx"], seed=1,
                           energy_type=args.energy_type, optimizer=args.optimizer)
    for _ in range(args.count_at):
        eng.step()
    torch.cuda.synchronize()
    calls = eng.stage_calls()
    torch.cuda.profiler.start()
    for st in stages_of(args.energy_type):
        _lib.check(calls[st](), lib)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    eng.close()


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
    ap.add_argument("--energy-type", default="graspqp", choices=["graspqp", "dexgrasp", "tdg"],
                    help="scripts/fit.py --energy_type (SURVEY 8(f3)); the headline metric is quoted on graspqp")
    ap.add_argument("--optimizer", default="mala_star", choices=["mala_star", "dexgraspnet"],
                    help="scripts/fit.py --optimizer (SURVEY 8(f3))")
    ap.add_argument("--no-graph", action="store_true", help="do not replay the energy evaluation from a CUDA graph")
    ap.add_argument("--no-overlap", action="store_true", help="run the penetration chain on the main stream too")
    ap.add_argument("--global-rms", action="store_true",
                    help="N>1: all-reduce the batch-wide MALA*/qpth statistics every iteration (shard.GlobalStats); the "
                         "default is the reference's practice of one independent process per object group")
    ap.add_argument("--shard-global", action="store_true",
                    help="N>1: deal the config's GLOBAL object list (config3: 64 meshes, config4: 256, config5: 32) "
                         "round-robin over the ranks instead of replicating the per-GPU list (strong scaling)")
    ap.add_argument("--cpu-sample-batch", type=int, default=16)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--sustain", type=int, default=2000,
                    help="N=1: iterations of the sustained block (windows of 250, outlier resets every --reset-epochs as "
                         "in scripts/fit.py:408-422); 0 = skip")
    ap.add_argument("--reset-epochs", type=int, default=600)
    ap.add_argument("--no-counters", action="store_true", help="skip the ncu subprocess (executed-instruction counters)")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip baseline G (the reference's GPU path)")
    ap.add_argument("--no-api-loop", action="store_true", help="skip timing the reference-shaped Python loop")
    ap.add_argument("--count-mode", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--count-at", type=int, default=25, help=argparse.SUPPRESS)
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
    w = workload.build(args.workload, rank=rank, batch_each=args.batch, n_objects=args.objects,
                       world=world if args.shard_global else None)
    B, n_c, O = w["batch_each"], w["n_contact"], len(w["face_verts"])
    Bt = B * O
    t_setup = time.perf_counter()
    eng = engine.FitEngine(lib, be, w["pack"], w["face_verts"], w["surface"], B, w["pose"], w["idx"],
                           qp_batch_stop=not args.no_qp_batch_stop, seed=1 + rank,
                           collectives=shard.GlobalStats() if (world > 1 and args.global_rms) else None,
                           overlap=not args.no_overlap, energy_type=args.energy_type, optimizer=args.optimizer)
    graphed = (not args.no_graph) and eng.capture()
    qp_batch_stop, overlap = eng.qp_batch_stop, eng.overlap
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
    # (with per-step collectives every rank must run the SAME number of iterations, and this loop's trip count depends on
    # each rank's nvidia-smi timing: a rank-local stage stands in for the load there — r2k: a full step here dead-locked
    # --global-rms on a collective sequence mismatch until NCCL's 10-minute watchdog)
    load_fn = eng.step if eng.collectives is None else eng.stage_calls()["hand_distance"]
    while sampler.proc is not None and sampler.count_since(t_timed) < 2 and time.perf_counter() - t_load < 3.0:
        flush.zero_()
        load_fn()
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
    import ctypes

    # ---- every stage alone on the current state (CUDA events on the launching stream, L2 flushed before each) ----
    def time_stages():
        out = {}
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
            out[name] = float(np.median(ts))
        return out

    stage_ms = time_stages()
    iteration_of_value = eng.iteration

    # ---- FP32 / FP64 FMA peaks, measured live (roofline denominators of the SIMT stages) ----
    def fma_peak(probe, dtype):
        sink = be.empty((4,), dtype)
        nf = ctypes.c_uint64(0)
        probe(2000, be.ptr(sink), ctypes.byref(nf), be.stream)
        torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        probe(32000, be.ptr(sink), ctypes.byref(nf), be.stream)
        f1.record()
        torch.cuda.synchronize()
        return 2.0 * nf.value / (f0.elapsed_time(f1) * 1e-3) / 1e12
    fp32_peak_tflops = fma_peak(lib.gqp_fma_peak_probe, np.float32)
    fp64_peak_tflops = fma_peak(lib.gqp_dfma_peak_probe, np.float64)

    # ---- sustained block: the fit carried on for --sustain iterations with the reference's outlier resets ----
    sustained = None
    if world == 1 and args.sustain > 0:
        from graspqp_b200 import initializations
        eng.enable_device_init([initializations.convex_hull(fv.reshape(-1, 3)) for fv in w["face_verts"]])
        i_pen = engine.TERMS.index("E_pen")
        win, windows, n_resets, n_reset_rows, reset_ms = 250, [], 0, 0, []
        n_total = args.sustain
        evs2 = []
        for it in range(1, n_total + 1):
            # scripts/fit.py:408: every reset_epochs iterations, not within the last 2*reset_epochs of the fit
            reset = bool(args.reset_epochs) and (eng.iteration + 1) % args.reset_epochs == 0 \
                and (eng.iteration + 1) < n_total - 2 * args.reset_epochs
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if reset:  # initialize_convex_hull on the device (3 launches) + gqp_mala_reset, INSIDE the timed step
                eng.reset_step(z_score_threshold=1.0)
            else:
                eng.step()
            e1.record()
            evs2.append((e0, e1))
            if reset:
                n_resets += 1
                n_reset_rows += int(eng.reset_count.item())
                reset_ms.append(e0.elapsed_time(e1))
            if it % win == 0 or it == n_total:
                torch.cuda.synchronize()
                ms = [a.elapsed_time(b) for a, b in evs2]
                evs2 = []
                cnt = eng.pen_counters()
                eng.assert_ok()
                windows.append({"iterations": [eng.iteration - len(ms) + 1, eng.iteration],
                                "ms_per_step": round(float(np.mean(ms)), 4), "ms_max": round(float(np.max(ms)), 4),
                                "mean_E_pen": round(float(eng.losses[i_pen].mean().item()), 5),
                                "mean_energy": round(float(eng.energy.mean().item()), 3),
                                "searched_samples_per_candidate": round(cnt[0] / Bt, 2),
                                "accept_rate": round(float(eng.accept.float().mean().item()), 3)})
        all_ms = sum(wd["ms_per_step"] * (wd["iterations"][1] - wd["iterations"][0] + 1) for wd in windows)
        sustained = {"iterations": n_total, "value": Bt * n_total / (all_ms * 1e-3), "unit": UNIT,
                     "ms_per_step": all_ms / n_total, "windows": windows, "resets": n_resets,
                     "reset_rows": n_reset_rows, "reset_epochs": args.reset_epochs,
                     "reset_init": "device (gqp_init_convex_hull: hull pool, FPS, poses) + gqp_mala_reset, inside the timed step",
                     "reset_step_ms": [round(v, 3) for v in reset_ms],
                     "value_window": f"the headline `value` times iterations {iteration_of_value - K + 1}-"
                                     f"{iteration_of_value} after init (K steps after W warm-up)",
                     "stage_ms_at_end": {k: round(v, 4) for k, v in time_stages().items()}}
    iteration_of_counters = eng.iteration

    # ---- max over ranks, gather ----
    ms_pen = stage_ms["hand_distance"]
    t = torch.tensor([ms_total, ms_e2e, ms_pen], dtype=torch.float64, device=dev)
    Bt_total, O_total = Bt * world, O * world
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot = torch.tensor([Bt, O], dtype=torch.int64, device=dev)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        Bt_total, O_total = int(tot[0]), int(tot[1])
        # final grasps of all ranks in global object order (the only data exchange of the path by default)
        shard.gather_results(eng.pose, eng.energy, eng.idx, O_total, B, rank, world)
    ms_total, ms_e2e, ms_pen = [float(v) for v in t.tolist()]
    final_energy = float(eng.energy.float().mean().item())
    eng.close()
    del eng
    torch.cuda.empty_cache()
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        f_hand = int(w["pack"]["face_offset"][-1])
        n_pts, L = w["surface"].shape[1], len(w["pack"]["mesh_frame"])
        # ---- executed-instruction counters of the three heavy stages, by an ncu subprocess of this run ----
        counters, counters_note = (None, "skipped") if (world > 1 or args.no_counters) else \
            collect_counters(args, iteration_of_counters)
        st_ms = sustained["stage_ms_at_end"] if sustained else stage_ms   # times at the state the counters were taken at
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        kernels = {}
        stages = stages_of(args.energy_type)
        for st in stages:
            row = {"ms": round(st_ms[st], 4)}
            if counters:
                c = counters[st]
                g = lambda k: c.get(k, 0.0)
                f32 = 2 * g(COUNTER_METRICS[0]) + g(COUNTER_METRICS[1]) + g(COUNTER_METRICS[2])
                f64 = 2 * g(COUNTER_METRICS[3]) + g(COUNTER_METRICS[4]) + g(COUNTER_METRICS[5])
                sec = st_ms[st] * 1e-3
                row.update({"fp32_tflops": f32 / sec / 1e12, "fp32_frac": f32 / sec / 1e12 / fp32_peak_tflops,
                            "fp64_tflops": f64 / sec / 1e12, "fp64_frac": f64 / sec / 1e12 / fp64_peak_tflops,
                            # every executed thread-instruction against the lane-issue roof (4 SMSPs x 32 lanes / SM / clk)
                            "lane_issue_frac": g("smsp__thread_inst_executed.sum") / (sec * sm_hz * sms * 128),
                            "lanes_per_inst": g("smsp__thread_inst_executed.sum") / max(g("smsp__inst_executed.sum"), 1.0),
                            "dram_bytes": g("dram__bytes_read.sum") + g("dram__bytes_write.sum")})
            kernels[st] = row
        top = max(stages, key=lambda st: st_ms[st])
        fp = "fp64" if top == "fc_solve" else "fp32"
        alg_bytes = {"hand_distance": Bt * (L * 48 + 4 + 4 * ((n_pts + 31) // 32)) + O * n_pts * 12,
                     "object_distance": Bt * n_c * (12 + 36), "fc_solve": Bt * n_c * 24 + Bt * 12 * 3 * 4 * n_c * 8,
                     "span_energy": Bt * (n_c * 76 + 8)}[top]
        roof = {"bound": fp, "kernel": f"{top} ({' + '.join(k.split('<')[0] for k in STAGE_KERNELS[top][:5] if '(int)' not in k)}): "
                                       "the longest stage of an iteration at the sustained state",
                "achieved": kernels[top].get(fp + "_tflops"), "peak": fp32_peak_tflops if fp == "fp32" else fp64_peak_tflops,
                "unit": "TFLOP/s", "frac": kernels[top].get(fp + "_frac"),
                "lane_issue_frac": kernels[top].get("lane_issue_frac"), "traffic": kernels[top].get("dram_bytes"),
                "ms_per_launch": st_ms[top], "counters": counters_note,
                "note": "achieved = EXECUTED FP flop (ncu: 2*ffma + fmul + fadd per launch group, predicated-on threads) / "
                        "CUDA-event time of the un-profiled launch; peak = FMA probe measured live in this run; "
                        "lane_issue_frac = all executed thread-instructions / (SMs x 128 lanes x SM clock x time), the "
                        "roof of a traversal kernel whose work is compares, shuffles and address arithmetic as much as "
                        "FP; traffic = dram bytes of the same ncu pass (cold caches)",
                "kernels": kernels,
                "bfe": {"flops_per_step": 80.0 * n_pts * f_hand * Bt, "speedup_vs_bruteforce_at_peak":
                        80.0 * n_pts * f_hand * Bt / (ms_pen * 1e-3) / 1e12 / fp32_peak_tflops,
                        "note": "SURVEY 8(d)'s brute-force-equivalent work of H4 (80 flop x samples x hand faces x "
                                "candidates) over the stage time, against the FMA peak: how much exact culling saves, "
                                "NOT a roofline fraction"},
                "hbm": {"achieved": alg_bytes / (st_ms[top] * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": alg_bytes / (st_ms[top] * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes": alg_bytes,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
        line = {
            "metric": METRIC, "value": Bt_total * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True,
            "scaling": "strong" if (args.shard_global and world > 1) else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"energy_type": args.energy_type, "optimizer": args.optimizer,
                       "workload": f"{args.workload}: {w['hand']}, {O} meshes/GPU ({','.join(w['kinds'])}; "
                                   f"{w['faces_per_mesh']} tris), batch {B}/object, n_contact {n_c}, 2500 surface samples",
                       "candidates_per_gpu": Bt, "objects_total": O_total, "candidates_total": Bt_total,
                       "sharding": ("global object list dealt round-robin over the ranks (shard.objects_of_rank)"
                                    if args.shard_global else "per-GPU object list replicated per rank"), "l2": "flushed between timed steps (256 MiB write)",
                       "qp_batch_stop": qp_batch_stop, "streams": 2 if overlap else 1, "cuda_graph": bool(graphed),
                       "collectives": ("global rms + qp exit words all-reduced per step" if args.global_rms else
                                       "none on the data path; final all-gather of grasps") if world > 1 else "n/a",
                       "noise": "philox on device (value) / injected from host (e2e)",
                       "e2e_pipeline": "double-buffered, copies on a second stream: upload k+1 and enqueue k+1 before reading back k",
                       "setup_s": round(t_setup, 2), "final_mean_energy": final_energy},
            "e2e": {"value": Bt_total * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "clocks": clocks,
            "stage_ms": {k: round(v, 4) for k, v in stage_ms.items()},
            "fp_peaks": {"fp32_tflops": fp32_peak_tflops, "fp64_tflops": fp64_peak_tflops,
                         "nominal_fp32_tflops": sms * 128 * 2 * sm_hz / 1e12},
            "roofline": roof,
        }
        if sustained:
            line["sustained"] = sustained
        if world == 1 and not args.no_api_loop:
            line["api_loop"] = api_loop_run(args, w, dev)
        if world == 1 and not args.no_gpu_baseline:
            line["gpu_baseline"] = gpu_baseline_run(args)
        if world == 1 and not args.skip_cpu_baseline:
            cb, _ = cpu_reference_run(args, 100, 1, args.cpu_sample_batch)   # ~15 s of host work
            line["cpu_baseline"] = cb
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def gpu_baseline_run(args):
    """Baseline G: the reference's GPU path on this GPU, in a subprocess (baseline/gpu_reference.py: the
    reference's MalaStar / calculate_energy / metric code + its TorchSDF kernel built for sm_100; qpth and the model
    classes substituted as stated there).  Bounded: 2 warm-up + 3 timed iterations of the same workload."""
    script = os.path.join(ROOT, "baseline", "gpu_reference.py")
    cmd = [sys.executable, script, "--workload", args.workload, "--steps", "3", "--warmup", "2"]
    if args.batch:
        cmd += ["--batch", str(args.batch)]
    if args.objects:
        cmd += ["--objects", str(args.objects)]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        last = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        return json.loads(last[-1]) if last else {"unavailable": f"rc {r.returncode}: {r.stderr[-200:]!r}"}
    except (subprocess.TimeoutExpired, OSError, ValueError) as e:
        return {"unavailable": f"{type(e).__name__}"}


def api_loop_run(args, w, dev, iters=10, warmup=3):
    """The reference-shaped Python loop (HandModel / ObjectModel / energy callable / MalaStar + autograd, the body of
    scripts/fit.py:399-458 written as the reference writes it) on the same workload: what a user gets who swaps the
    imports and keeps their loop, next to `value` (the fused FitEngine)."""
    import torch
    import graspqp_b200 as g
    from graspqp_b200 import engine
    B, n_c, O = w["batch_each"], w["n_contact"], len(w["face_verts"])
    hand = g.HandModel(w["hand"], dev)
    obj = g.ObjectModel(batch_size_each=B, num_samples=w["surface"].shape[1], device=dev)
    obj.initialize(list(w["kinds"]), w["face_verts"], w["surface"])
    F = g.GraspSpanMetricFactory
    energy_fnc = F.create(F.MetricType.GRASPQP, solver_kwargs={"friction": 0.2, "max_limit": 20.0, "n_cone_vecs": 4})
    hand.set_parameters(torch.tensor(w["pose"], device=dev).requires_grad_(True), torch.tensor(w["idx"], device=dev))
    opt = g.MalaStar(hand, switch_possibility=0.4, device=dev, batch_size=B)
    wts = engine.DEFAULT_WEIGHTS

    def total(losses):
        e = 0
        for k, v in losses.items():
            e = e + wts[k] * v
        return e
    energy = total(g.calculate_energy(hand, obj, energy_fnc=energy_fnc, method="gendexgrasp", svd_gain=0.1))
    energy.sum().backward()
    opt.zero_grad()
    energy = energy.detach()

    def iterate():
        nonlocal energy
        opt.try_step()
        Eb = energy.view(-1, B)
        z = ((Eb - Eb.mean(-1).unsqueeze(-1)) / Eb.std(-1).unsqueeze(-1)).view(-1)
        opt.zero_grad()
        new_energy = total(g.calculate_energy(hand, obj, energy_fnc=energy_fnc, method="gendexgrasp", svd_gain=0.1))
        new_energy.sum().backward()
        with torch.no_grad():
            accept, _ = opt.accept_step(energy, new_energy, None, z, 1.0)
            energy[accept] = new_energy.detach()[accept]
    for _ in range(warmup):
        iterate()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        iterate()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / iters
    return {"value": O * B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "iterations": iters,
            "what": "graspqp_b200.HandModel / ObjectModel / calculate_energy / MalaStar driven by torch autograd "
                    "(ops.py bridges), un-fused: fresh tensors per op, per-candidate surface samples, un-culled H4"}


if __name__ == "__main__":
    a = parse()
    if a.count_mode:
        run_count_mode(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
