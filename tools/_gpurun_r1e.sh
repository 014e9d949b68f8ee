set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python bench.py --workload config5 --skip-cpu-baseline > gpurun_out/r1g_bench_config5.json 2> gpurun_out/r1g_bench_config5.err
timeout 600 python bench.py --workload config4_robotiq2 --objects 8 --skip-cpu-baseline > gpurun_out/r1g_bench_config4_robotiq2.json 2> gpurun_out/r1g_bench_config4_robotiq2.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r1g_bench_reference_arm.json 2> gpurun_out/r1g_bench_reference_arm.err
tail -n 2 gpurun_out/r1g_*.err
