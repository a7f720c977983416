#!/bin/bash
# pool pipeline (flags=512) of the working tree and of every variant library vs the persistent kernel:  bash tools/pool_ab.sh TAG
TAG=${1:-pab}; O=gpurun_out/$TAG; mkdir -p $O
timeout 300 python tools/pool_check.py 2>&1 | tail -2 | tee -a $O/times.txt
for w in "C2 --spp 64" "C2 --spp 256" "C2classic --spp 64" "C4 --spp 16" "C3 --spp 8"; do
  timeout 200 python tools/prof_run.py $w --reps 3 2>&1 | tail -1 | tee -a $O/times.txt
  timeout 200 python tools/prof_run.py $w --reps 3 --flags 512 2>&1 | tail -1 | tee -a $O/times.txt
  for lib in raytracingoneweekend_b200/_variants/libom_*.so; do
    [ -e "$lib" ] || continue
    echo -n "$(basename $lib) " | tee -a $O/times.txt
    OM_LIB_PATH=$lib timeout 200 python tools/prof_run.py $w --reps 3 --flags 512 2>&1 | tail -1 | tee -a $O/times.txt
  done
done
