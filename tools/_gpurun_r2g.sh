#!/usr/bin/env bash
# launch lists (cold-cache, serialised: compare shares) of config 4 / Panda and config 3 at 8 objects per GPU
mkdir -p gpurun_out
F="--sustain 0 --no-counters --no-gpu-baseline --no-api-loop --skip-cpu-baseline --no-graph"
for wl in config4 config3; do
python bench.py --workload $wl --objects 8 --steps 2 --warmup 3 $F > gpurun_out/r2g_plain_$wl.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 150 --csv --log-file gpurun_out/r2g_launches_$wl.csv \
  python bench.py --workload $wl --objects 8 --steps 2 --warmup 3 $F > gpurun_out/r2g_ncu_$wl.log 2>&1
done
ls -la gpurun_out/r2g_*
