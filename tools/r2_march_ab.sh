#!/bin/bash
TAG=${1:-mab}; O=gpurun_out/$TAG; mkdir -p $O
python -m pytest tests -m gpu -x -q -k "marched or mixed or dyn or adaptive or smoke or full_resolution" > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/times.txt; tail -2 $O/pytest.log
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  echo "== $lib" | tee -a $O/times.txt
  for w in "C3 --spp 8" "C2 --spp 64" "C4 --spp 16"; do
    if [ "$lib" = default ]; then timeout 300 python tools/prof_run.py $w --reps 3 2>&1 | tail -1 | tee -a $O/times.txt
    else OM_LIB_PATH=$lib timeout 300 python tools/prof_run.py $w --reps 3 2>&1 | tail -1 | tee -a $O/times.txt; fi
  done
done
