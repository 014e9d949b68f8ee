#!/usr/bin/env bash
# round 2, first GPU pass: parity tests, baseline G, a bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 600 python baseline/gpu_reference.py --workload config2 --steps 5 --warmup 2 > gpurun_out/r2a_G_config2.json 2> gpurun_out/r2a_G_config2.err; tail -c 600 gpurun_out/r2a_G_config2.json; tail -3 gpurun_out/r2a_G_config2.err
timeout 600 python baseline/gpu_reference.py --workload headline --objects 1 --steps 5 --warmup 2 > gpurun_out/r2a_G_headline1.json 2> gpurun_out/r2a_G_headline1.err; tail -c 600 gpurun_out/r2a_G_headline1.json
python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench_config2.json 2> gpurun_out/r2a_bench.err; tail -c 1500 gpurun_out/r2a_bench_config2.json
