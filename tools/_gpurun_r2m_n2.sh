#!/usr/bin/env bash
mkdir -p gpurun_out
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus 2 --steps 10 --warmup 3 $F --workload config2 --global-rms > gpurun_out/r2m_config2_global_rms_n2.json 2> gpurun_out/r2m.err
echo "rc=$?"; grep -E "Error|error|Traceback" gpurun_out/r2m.err | head -5
python - <<'PY'
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2m_config2_global_rms_n2.json") if l.startswith("{")][-1])
    print("global-rms N=2: %.2fM grasp-steps/s, %.3f ms/step, e2e %.2fM, %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["config"]["collectives"]))
except Exception as e: print("failed", e)
PY
