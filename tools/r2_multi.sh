#!/bin/bash
# N-GPU validation:  gpurun --gpus N -- 'bash tools/r2_multi.sh TAG N'
TAG=${1:-m}; N=${2:-2}; O=gpurun_out/$TAG; mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; }
run --steps 5 --warmup 3 > $O/bench_C2_${N}gpu.json 2> $O/bench_C2_${N}gpu.err; echo "C2 rc=$?" | tee -a $O/rc.txt
run --adaptive --spp 256 --steps 3 --warmup 1 > $O/bench_C2_adaptive_${N}gpu.json 2> $O/bench_C2_adaptive_${N}gpu.err; echo "adaptive rc=$?" | tee -a $O/rc.txt
python -m pytest tests -m gpu -x -q -k "dist or multi" > $O/pytest.log 2>&1; tail -2 $O/pytest.log
cat $O/bench_C2_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d.get('phase_ms'), d.get('merge_check'))"
cat $O/bench_C2_adaptive_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d.get('merge_check'))"
for f in $O/*.err; do tail -n 3 $f; done
