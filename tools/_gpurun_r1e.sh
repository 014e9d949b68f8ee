set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15) > gpurun_out/r1f_pytest.log
(timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3) > gpurun_out/r1f_smoke.log
timeout 600 python bench.py > gpurun_out/r1f_bench_config2.json 2> gpurun_out/r1f_bench_config2.err
timeout 600 python bench.py --workload config3 --skip-cpu-baseline > gpurun_out/r1f_bench_config3.json 2> gpurun_out/r1f_bench_config3.err
timeout 600 python bench.py --workload headline --skip-cpu-baseline > gpurun_out/r1f_bench_headline.json 2> gpurun_out/r1f_bench_headline.err
timeout 600 python bench.py --workload config4 --objects 8 --skip-cpu-baseline > gpurun_out/r1f_bench_config4_panda.json 2> gpurun_out/r1f_bench_config4_panda.err
tail -5 gpurun_out/r1f_pytest.log; cat gpurun_out/r1f_smoke.log; tail -n 2 gpurun_out/r1f_*.err
for f in config2 config3 headline config4_panda; do python -c "
import json; d=json.loads(open('gpurun_out/r1f_bench_$f.json').read().strip().splitlines()[-1]); print('$f', d['value'], d['ms_per_step'], d['e2e']['value'], d['stage_ms'])"; done
