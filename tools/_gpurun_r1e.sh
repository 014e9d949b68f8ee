set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r1e_pytest.log
timeout 600 python bench.py > gpurun_out/r1e_bench_config2.json 2> gpurun_out/r1e_bench_config2.err
timeout 600 python bench.py --workload config4_robotiq2 --objects 8 --skip-cpu-baseline > gpurun_out/r1e_bench_config4_robotiq2.json 2> gpurun_out/r1e_bench_config4_robotiq2.err
timeout 600 python bench.py --workload config4 --objects 8 --skip-cpu-baseline > gpurun_out/r1e_bench_config4_panda.json 2> gpurun_out/r1e_bench_config4_panda.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r1e_bench_reference_arm.json 2> gpurun_out/r1e_bench_reference_arm.err
tail -3 gpurun_out/r1e_pytest.log; tail -2 gpurun_out/*.err; cat gpurun_out/r1e_bench_config4_robotiq2.json gpurun_out/r1e_bench_config4_panda.json gpurun_out/r1e_bench_reference_arm.json
