#!/usr/bin/env bash
# round 2, last GPU call: the (f3)/(f4) GPU tests first (new code), then the whole -m gpu suite, then a short bench line
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_y_next_rows_gpu.py tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/r2n_pytest_new.log 2>&1
echo "new rc=$?" | tee -a gpurun_out/r2n_pytest_new.log
tail -5 gpurun_out/r2n_pytest_new.log
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_pytest_all.log 2>&1
echo "all rc=$?" | tee -a gpurun_out/r2n_pytest_all.log
tail -3 gpurun_out/r2n_pytest_all.log
true

