#!/bin/bash
# Round check on a B200 box: GPU tests, smoke, the bench lines, the ncu launch list and one full capture of k_render.
# Usage: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh TAG'
TAG=${1:-head}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc.txt
python bench.py > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench rc=$?" | tee -a $O/rc.txt
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_launch.log 2>&1; echo "ncu-launch rc=$?" | tee -a $O/rc.txt
ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o $O/k_render_c2 -f python tools/prof_run.py C2 --spp 16 --reps 1 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
tail -3 $O/pytest_gpu.log; cat $O/bench_c2.json
