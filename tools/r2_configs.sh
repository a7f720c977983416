#!/bin/bash
# every config with cpu_baseline + ncu of the dominant kernels of C3 and C5:  bash tools/r2_configs.sh TAG
TAG=${1:-cfg}; O=gpurun_out/$TAG; mkdir -p $O
for w in C1 C2classic C3 C4 C5; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > $O/bench_${w}_1gpu.json 2> $O/bench_${w}.err; echo "$w rc=$?" | tee -a $O/rc.txt
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o $O/k_render_C3 -f python tools/prof_run.py C3 --spp 4 --reps 1 > $O/ncu_C3.log 2>&1; echo "ncu C3 rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_extend --launch-skip 40 -c 1 -o $O/k_wf_extend_C5 -f python tools/prof_run.py C5 --spp 2 --reps 1 > $O/ncu_C5e.log 2>&1; echo "ncu C5 extend rc=$?" | tee -a $O/rc.txt
timeout 900 ncu --set full --clock-control none -k regex:k_wf_shade --launch-skip 40 -c 1 -o $O/k_wf_shade_C5 -f python tools/prof_run.py C5 --spp 2 --reps 1 > $O/ncu_C5s.log 2>&1; echo "ncu C5 shade rc=$?" | tee -a $O/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_C2.csv python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_launch.log 2>&1; echo "ncu-launch rc=$?" | tee -a $O/rc.txt
