set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/r1e_pytest.log
(timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5) > gpurun_out/r1e_smoke.log
timeout 600 python bench.py > gpurun_out/r1e_bench_config2.json 2> gpurun_out/r1e_bench_config2.err
tail -3 gpurun_out/r1e_pytest.log; cat gpurun_out/r1e_smoke.log; cat gpurun_out/r1e_bench_config2.json
