#!/usr/bin/env bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -p no:cacheprovider -k "benched or culling or hand_distance" > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
grep -E "^FAILED|passed|failed" gpurun_out/r2f_pytest.log | tail -8
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
for wl in config2 config4 config4_robotiq2 config3 config5; do
  python bench.py --workload $wl --steps 10 --warmup 3 $F > gpurun_out/r2f_bench_${wl}_n1.json 2>> gpurun_out/r2f.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2f_bench_${wl}_n1.json"))
print("$wl: %.2fM grasp-steps/s, %.3f ms/step, e2e %.2fM, Bt %d, stages %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["config"]["candidates_per_gpu"], d["stage_ms"]))
PY
done
