#!/bin/bash
O=gpurun_out/s3; mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
for spp in 16 256 1024; do python tools/counters.py C2 --spp $spp 2>&1 | tail -2 | tee -a $O/counters.txt; done
python tools/counters.py C4 --spp 64 2>&1 | tail -2 | tee -a $O/counters.txt
ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o $O/k_render_C2_256 -f python tools/prof_run.py C2 --spp 256 --reps 1 --flags 0 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
