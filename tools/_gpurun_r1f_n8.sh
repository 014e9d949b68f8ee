set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python bench.py --skip-cpu-baseline > gpurun_out/r1g_bench_config2_n1.json 2> gpurun_out/r1g_bench_config2_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r1g_bench_config2_n$N.json 2> gpurun_out/r1g_bench_config2_n$N.err
tail -n 3 gpurun_out/r1g_bench_config2_n*.err; cut -c1-200 gpurun_out/r1g_bench_config2_n*.json
