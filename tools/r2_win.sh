#!/bin/bash
# A/B of the working tree against _variants/libom_*.so at the spp that matter (16, 128, 1024) plus the lane occupancy, then the GPU tests.
#   gpurun --timeout 1500 -- 'bash tools/r2_win.sh TAG'
TAG=${1:-win}; O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
run() { if [ "$1" = default ]; then timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1; else OM_LIB_PATH=$1 timeout 300 python tools/prof_run.py $2 --spp $3 --reps 3 2>&1 | tail -1; fi; }
for lib in default raytracingoneweekend_b200/_variants/libom_*.so; do
  [ -e "$lib" ] || [ "$lib" = default ] || continue
  echo "== $lib" | tee -a $O/times.txt
  for w in "C2 16" "C2 128" "C2 1024" "C2classic 64" "C4 32" "C1 10" "C3 8"; do run $lib $w | tee -a $O/times.txt; done
done
for s in 16 256; do python tools/counters.py C2 --spp $s 2>&1 | tail -2 | tee -a $O/times.txt; done
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/times.txt; tail -5 $O/pytest.log
