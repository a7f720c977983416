#!/bin/bash
# Every built variant (and the default library) x the given render flags:  bash tools/variants_flags_run.sh TAG WORKLOAD SPP FLAGS...
TAG=$1; WL=$2; SPP=$3; shift 3
O=gpurun_out/$TAG; mkdir -p $O
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  for f in "$@"; do
    echo "$lib flags=$f" | tee -a $O/times.txt
    if [ "$lib" = default ]; then timeout 300 python tools/prof_run.py $WL --spp $SPP --reps 3 --flags $f 2>&1 | tail -2 | tee -a $O/times.txt
    else OM_LIB_PATH=$lib timeout 300 python tools/prof_run.py $WL --spp $SPP --reps 3 --flags $f 2>&1 | tail -2 | tee -a $O/times.txt; fi
  done
done
