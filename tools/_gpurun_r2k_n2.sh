#!/usr/bin/env bash
mkdir -p gpurun_out
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$T --master-port 29521 bench.py --gpus 2 --steps 10 --warmup 3 $F --workload config2 --global-rms > gpurun_out/r2k_config2_global_rms_n2.json 2> gpurun_out/r2k_global_rms_n2.err
echo "global-rms rc=$?"; grep -v "^\*\|OMP_NUM" gpurun_out/r2k_global_rms_n2.err | tail -15; tail -c 300 gpurun_out/r2k_config2_global_rms_n2.json
$T --master-port 29522 bench.py --gpus 2 --steps 10 --warmup 3 $F --workload config2 > gpurun_out/r2k_config2_n2.json 2>> gpurun_out/r2k_n2.err
$T --master-port 29523 bench.py --gpus 2 --steps 10 --warmup 3 $F --workload config5 --shard-global > gpurun_out/r2k_config5_n2.json 2>> gpurun_out/r2k_n2.err
for f in config2_global_rms config2 config5; do python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2k_${f}_n2.json") if l.startswith("{")][-1])
    print("$f N=2: %.2fM grasp-steps/s, %.3f ms/step, e2e %.2fM, %d candidates, %s, %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["config"]["candidates_total"], d["scaling"], d["config"]["collectives"]))
except Exception as e: print("$f failed", e)
PY
done
