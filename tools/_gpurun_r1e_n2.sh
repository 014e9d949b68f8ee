set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r1e_bench_config2_n2.json 2> gpurun_out/r1e_bench_config2_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --global-rms > gpurun_out/r1e_bench_config2_n2_globalrms.json 2> gpurun_out/r1e_bench_config2_n2_globalrms.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 5 --warmup 1 > gpurun_out/r1e_bench_reference_n2.json 2> gpurun_out/r1e_bench_reference_n2.err
tail -n 4 gpurun_out/r1e_bench_config2_n2.err gpurun_out/r1e_bench_config2_n2_globalrms.err gpurun_out/r1e_bench_reference_n2.err
cat gpurun_out/r1e_bench_config2_n2.json gpurun_out/r1e_bench_config2_n2_globalrms.json gpurun_out/r1e_bench_reference_n2.json
