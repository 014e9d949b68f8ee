#!/usr/bin/env bash
# BASELINE.json's multi-GPU configs, sharded by object over the 8 GPUs of one box (global object lists)
mkdir -p gpurun_out
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
N=${1:-8}
run() { # name, extra args
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 $F "${@:2}" > gpurun_out/r2h_$1_n$N.json 2>> gpurun_out/r2h_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2h_$1_n$N.json") if l.startswith("{")][-1])
    print("$1 N=$N: %.2fM grasp-steps/s, %.3f ms/step, e2e %.2fM, %d candidates on %d objects, %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["config"]["candidates_total"], d["config"]["objects_total"], d["scaling"]))
except Exception as e:
    print("$1 N=$N failed:", e)
PY
}
if [ "$N" = "8" ]; then
run config3 --workload config3 --shard-global
run config4_panda --workload config4 --shard-global
run config4_robotiq2 --workload config4_robotiq2 --shard-global
run config2_global_rms --workload config2 --global-rms
fi
run config5 --workload config5 --shard-global
tail -5 gpurun_out/r2h_n$N.err
