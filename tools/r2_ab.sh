#!/bin/bash
# A/B session on a B200 box: GPU tests of the working tree, then k_render timings of the working tree and of every
# library under raytracingoneweekend_b200/_variants/ on a few workloads.
#   gpurun --timeout 900 -- 'bash tools/r2_ab.sh TAG [notest]'
TAG=${1:-ab}; O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
if [ "$2" != "notest" ]; then ( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -5 $O/pytest_gpu.log; fi
run() {  # lib workload spp
  if [ "$1" = default ]; then timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1
  else OM_LIB_PATH=$1 timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1; fi
}
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  echo "== $lib" | tee -a $O/times.txt
  for w in "C2 16" "C2 64" "C2 256" "C2classic 64" "C4 16" "C1 10" "C3 8"; do run $lib $w | tee -a $O/times.txt; done
done
python tools/counters.py C2 --spp 16 2>&1 | tail -2 | tee -a $O/times.txt
