#!/bin/bash
# 1-GPU validation of bench.py + fresh ncu of k_render at HEAD:  bash tools/r2_bench1.sh TAG
TAG=${1:-b1}; O=gpurun_out/$TAG; mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
python bench.py > $O/bench_C2.json 2> $O/bench_C2.err; echo "bench C2 rc=$?" | tee -a $O/rc.txt
python bench.py --workload C1 --steps 5 > $O/bench_C1.json 2> $O/bench_C1.err; echo "bench C1 rc=$?" | tee -a $O/rc.txt
python bench.py --adaptive --spp 256 --no-cpu > $O/bench_C2_adaptive.json 2> $O/bench_C2_adaptive.err; echo "bench adaptive rc=$?" | tee -a $O/rc.txt
python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/rc.txt
ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -o $O/k_render_C2 -f python tools/prof_run.py C2 --spp 16 --reps 1 > $O/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $O/rc.txt
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_bytes.sum --clock-control none -k regex:k_render --launch-skip 1 -c 1 -o $O/k_render_C2_1024_dram -f python tools/prof_run.py C2 --spp 1024 --reps 1 > $O/ncu_dram.log 2>&1; echo "ncu-dram rc=$?" | tee -a $O/rc.txt
cat $O/bench_C2.json | cut -c1-400
