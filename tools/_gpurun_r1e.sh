set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r1f_pytest.log
timeout 600 python bench.py --workload config4 --objects 8 --skip-cpu-baseline > gpurun_out/r1f_bench_config4_panda.json 2> gpurun_out/r1f_bench_config4_panda.err
tail -3 gpurun_out/r1f_pytest.log; tail -n 2 gpurun_out/r1f_bench_config4_panda.err; cat gpurun_out/r1f_bench_config4_panda.json
