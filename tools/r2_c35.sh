#!/bin/bash
# C3 / C5 A/B of the working tree against _variants/libom_*.so, then the GPU tests
TAG=${1:-c35}; O=gpurun_out/$TAG; mkdir -p $O
run() { if [ "$1" = default ]; then timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1; else OM_LIB_PATH=$1 timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1; fi; }
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  echo "== $lib" | tee -a $O/times.txt
  for w in "C3 8" "C3 32" "C5 4" "C2 64"; do run $lib $w | tee -a $O/times.txt; done
done
python tools/counters.py C5 --spp 4 --freeze 16 2>&1 | tail -1 | cut -c1-300 | tee -a $O/times.txt
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/times.txt; tail -3 $O/pytest.log
