#!/bin/bash
# One full ncu capture of k_render on a workload:  bash tools/r2_ncu.sh TAG [WORKLOAD] [SPP] [test]
TAG=${1:-ncu}; WL=${2:-C2}; SPP=${3:-16}; O=gpurun_out/$TAG; mkdir -p $O
if [ "$4" = "test" ]; then ( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log; fi
python tools/prof_run.py $WL --spp $SPP --reps 3 2>&1 | tail -1 | tee $O/time.txt
ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o $O/k_render_$WL -f python tools/prof_run.py $WL --spp $SPP --reps 1 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
