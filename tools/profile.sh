#!/bin/bash
# tools/profile.sh TAG [EVENTS] — the profiling recipe of B200_PROFILING.md for the bench workload, one GPU:
#   1. the plain run (must exit 0 before any ncu pass; its JSON line is the only bench value),
#   2. the launch list (gpu__time_duration per launch),
#   3. one --set full capture of the two hot kernels with source correlation.
# Outputs land in gpurun_out/ (scratch); summaries are copied to profiles/ by hand.
set -e
TAG=${1:-rXX}; EV=${2:-8000000}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --events $EV --steps 2 --warmup 3 > gpurun_out/plain_$TAG.log 2>&1
tail -1 gpurun_out/plain_$TAG.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --events $EV --steps 1 --warmup 1 > gpurun_out/ncu_launch_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k1_decode|k2_rules_bucket' -s 4 -c 2 -f -o gpurun_out/prof_$TAG \
    python bench.py --events $EV --steps 1 --warmup 1 > gpurun_out/ncu_full_$TAG.log 2>&1
ls -la gpurun_out/prof_$TAG.ncu-rep
