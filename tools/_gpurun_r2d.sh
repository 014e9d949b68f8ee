#!/usr/bin/env bash
# perf pass: fused FK kernels; residency of the persistent search kernel vs the critical chain
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 5 --sustain 750 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
for k in 8 6 4 3; do
  GQP_PEN_SEARCH_CTAS=$k $B > gpurun_out/r2d_bench_pen$k.json 2>> gpurun_out/r2d.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2d_bench_pen$k.json")); s=d["sustained"]
print("pen_search CTAs/SM $k: value %.3fM ms %.4f | sustained ms %.4f last window %.4f | stages %s" % (d["value"]/1e6, d["ms_per_step"], s["ms_per_step"], s["windows"][-1]["ms_per_step"], s["stage_ms_at_end"]))
PY
done
