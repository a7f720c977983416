#!/bin/bash
# Time k_render on a workload for the default library and every built variant:  bash tools/variants_run.sh TAG [WORKLOAD] [SPP]
TAG=${1:-var}; WL=${2:-C2}; SPP=${3:-64}
O=gpurun_out/$TAG; mkdir -p $O
echo "default" | tee -a $O/times.txt
timeout 120 python tools/prof_run.py $WL --spp $SPP --reps 4 2>&1 | tail -3 | tee -a $O/times.txt
for f in raytracingoneweekend_b200/_variants/libom_*.so; do
  echo "$f" | tee -a $O/times.txt
  OM_LIB_PATH=$f timeout 120 python tools/prof_run.py $WL --spp $SPP --reps 4 2>&1 | tail -3 | tee -a $O/times.txt
done
