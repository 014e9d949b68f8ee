set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/r1f_ncu_plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r1f_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --no-graph > gpurun_out/r1f_ncu_launches.log 2>&1
timeout 300 python bench.py --workload config4 --objects 8 --steps 2 --warmup 3 --skip-cpu-baseline > gpurun_out/r1f_ncu_plain4.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fc_dis_kernel|pen_search_kernel|object_distance_kernel' --launch-skip 9 -c 4 -o gpurun_out/r1f_config4_panda_full python bench.py --workload config4 --objects 8 --steps 2 --warmup 3 --skip-cpu-baseline --no-graph > gpurun_out/r1f_ncu_full4.log 2>&1
ls -la gpurun_out/
