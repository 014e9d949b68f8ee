#!/usr/bin/env bash
# final evidence of round 2: the no-alloc test alone, all GPU tests, the default bench line, ncu launch list + full capture
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r2i_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2i_pytest.log
grep -E "^FAILED|passed|failed" gpurun_out/r2i_pytest.log | tail -8
python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench_config2.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r2i_bench_config2.json; tail -3 gpurun_out/r2i_bench.err
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline"
python bench.py --steps 3 --warmup 3 $F > gpurun_out/r2i_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 90 -c 120 --csv --log-file gpurun_out/r2i_launches_bench_steps3.csv \
  python bench.py --steps 3 --warmup 3 $F > gpurun_out/r2i_ncu_launches.log 2>&1
python bench.py --steps 3 --warmup 3 $F > gpurun_out/r2i_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'object_distance_kernel|pen_search_kernel|pen_collect_kernel|fc_dis_kernel' -s 12 -c 5 \
  -o gpurun_out/r2i_prof python bench.py --steps 3 --warmup 3 $F > gpurun_out/r2i_ncu_full.log 2>&1
ls -la gpurun_out | tail -12
