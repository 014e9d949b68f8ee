#!/usr/bin/env bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
grep -E "^FAILED|passed|failed" gpurun_out/r2c_pytest.log | tail -12
python bench.py --steps 20 --warmup 5 > gpurun_out/r2c_bench_config2.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2c_bench_config2.json; tail -5 gpurun_out/r2c_bench.err
