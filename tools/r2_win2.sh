#!/bin/bash
# window vs tiles with and without the phase barriers
TAG=${1:-win2}; O=gpurun_out/$TAG; mkdir -p $O
run() { if [ "$1" = default ]; then timeout 300 python tools/prof_run.py $2 --spp $3 --flags $4 --reps 3 2>&1 | tail -1; else OM_LIB_PATH=$1 timeout 300 python tools/prof_run.py $2 --spp $3 --flags $4 --reps 3 2>&1 | tail -1; fi; }
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  echo "== $lib" | tee -a $O/times.txt
  for w in "C2 128 0" "C2 128 256" "C2 512 0" "C2 512 256" "C2classic 64 256"; do run $lib $w | tee -a $O/times.txt; done
done
