#!/bin/bash
# Time k_render on a workload with different render flags: bash tools/flags_run.sh TAG WORKLOAD SPP FLAGS...
TAG=$1; WL=$2; SPP=$3; shift 3
O=gpurun_out/$TAG; mkdir -p $O
for f in "$@"; do timeout 300 python tools/prof_run.py $WL --spp $SPP --reps 4 --flags $f 2>&1 | tail -3 | tee -a $O/times.txt; done
