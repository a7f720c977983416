#!/bin/bash
# N-GPU bench lines at HEAD (fixed spp; adaptive merged at the end and after every pass):  gpurun --gpus N -- 'bash tools/r2_scale2.sh TAG N'
TAG=${1:-sc}; N=${2:-8}; O=gpurun_out/$TAG; mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; }
run --steps 6 --warmup 3 > $O/bench_C2_${N}gpu.json 2> $O/bench_C2_${N}gpu.err; echo "C2 rc=$?" | tee -a $O/rc.txt
run --adaptive --spp 256 --steps 4 --warmup 2 --no-cpu > $O/bench_C2_adaptive_${N}gpu.json 2> $O/bench_C2_adaptive_${N}gpu.err; echo "adaptive rc=$?" | tee -a $O/rc.txt
run --adaptive --spp 256 --steps 4 --warmup 2 --no-cpu --refresh-every 1 > $O/bench_C2_adaptive_refresh1_${N}gpu.json 2> $O/bench_C2_adaptive_refresh1_${N}gpu.err; echo "adaptive refresh rc=$?" | tee -a $O/rc.txt
for f in $O/bench_C2_${N}gpu.json $O/bench_C2_adaptive_${N}gpu.json $O/bench_C2_adaptive_refresh1_${N}gpu.json; do python -c "
import json,sys
d=json.loads(open('$f').read()); print(round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), d.get('phase_ms'), (d.get('merge_check') or {}).get('ok'), (d.get('merge_check') or {}).get('bit_identical'))"; done
