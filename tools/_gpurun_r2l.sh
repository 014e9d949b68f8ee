#!/usr/bin/env bash
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/r2l_bench_config2.json 2> gpurun_out/r2l_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2l_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2l_bench_config2.json")); r=d["roofline"]
print("value %.3fM, sustained %.3fM, roofline frac %s lane %s traffic %s | %s" % (d["value"]/1e6, d["sustained"]["value"]/1e6, r["frac"], r["lane_issue_frac"], r["traffic"], r["counters"]))
PY
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2l_pytest.log
grep -E "^FAILED|passed|failed" gpurun_out/r2l_pytest.log | tail -6
