#!/bin/bash
# A/B of an environment switch:  bash tools/ab_env_run.sh TAG VAR WORKLOAD SPP [WORKLOAD SPP ...]
TAG=$1; VAR=$2; shift 2
O=gpurun_out/$TAG; mkdir -p $O
while [ $# -ge 2 ]; do
  echo "$1 $2 spp: default" | tee -a $O/times.txt; timeout 300 python tools/prof_run.py $1 --spp $2 --reps 3 2>&1 | tail -2 | tee -a $O/times.txt
  echo "$1 $2 spp: $VAR=1" | tee -a $O/times.txt; env $VAR=1 timeout 300 python tools/prof_run.py $1 --spp $2 --reps 3 2>&1 | tail -2 | tee -a $O/times.txt
  shift 2
done
