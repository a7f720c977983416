#!/bin/bash
# Copy a measurement pass (gpurun_out/TAG, made by tools/r2_bench1.sh + tools/r2_configs.sh) into profiles/ as r2_TAG_* and
# refresh profiles/ncu_current.json from its captures:  bash tools/r2_collect.sh v2d
TAG=$1; V=gpurun_out/$TAG; P=profiles/r2_$TAG
for k in "k_render_C2 c2_k_render" "k_render_C3 c3_k_render" "k_wf_extend_C5 c5_k_wf_extend" "k_wf_shade_C5 c5_k_wf_shade"; do set -- $k; [ -e $V/$1.ncu-rep ] && ncu -i $V/$1.ncu-rep --page details 2>/dev/null > ${P}_$2.txt; done
cp $V/launches_C2.csv ${P}_c2_launches.csv
ncu -i $V/k_render_C2_1024_dram.ncu-rep --page raw --csv 2>/dev/null > ${P}_c2_1024spp_dram.csv
for c in C1 C2classic C3 C4 C5; do cp $V/bench_${c}_1gpu.json ${P}_bench_${c}_1gpu.json; done
cp $V/bench_C2.json ${P}_bench_C2_1gpu.json; cp $V/bench_C2_adaptive.json ${P}_bench_C2_adaptive256_1gpu.json; cp $V/bench_ref.json ${P}_bench_reference_C2.json
python tools/ncu_to_json.py C2 $V/k_render_C2.ncu-rep --kernel "k_render<" --spp 16 --dram $V/k_render_C2_1024_dram.ncu-rep --dram-spp 1024 --profile-file ${P}_c2_k_render.txt > /dev/null
python tools/ncu_to_json.py C3 $V/k_render_C3.ncu-rep --kernel "k_render<" --spp 4 --profile-file ${P}_c3_k_render.txt > /dev/null
python tools/ncu_to_json.py C5 $V/k_wf_extend_C5.ncu-rep --kernel "k_wf_extend" --spp 2 --profile-file ${P}_c5_k_wf_extend.txt > /dev/null
python - <<PY
import json
d=json.load(open('profiles/ncu_current.json'))
for k in ('C2','C3','C5'):
    e=d[k]; print(k, e['kernel'], 'ms %.3f lanes %.2f issue %.1f%% fma %.1f%% l2hit %.1f%% long_sb %.2f barrier %.2f'%(e['time_ms'],e['threads_per_warp_instruction'],e['issue_active_pct'],e['pipe_fma_cycles_active_pct'],e['l2_hit_pct'],e['stall_long_scoreboard_per_issue'],e['stall_barrier_per_issue']), e.get('dram_bytes'))
PY
