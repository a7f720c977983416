#!/bin/bash
# instruction counts of k_render: working tree vs _variants/libom_tiles.so (C2, 64 spp)
TAG=${1:-win3}; O=gpurun_out/$TAG; mkdir -p $O
M=smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,smsp__inst_executed_op_global_ld.sum,sm__cycles_elapsed.max,smsp__issue_active.sum,smsp__inst_executed_op_branch.sum,sm__inst_executed_pipe_cbu.sum,smsp__warp_issue_stalled_barrier_per_warp_active.pct,smsp__warp_issue_stalled_no_instruction_per_warp_active.pct,smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct,smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct
ncu --metrics $M --clock-control none -k regex:k_render --csv --log-file $O/new.csv python tools/prof_run.py C2 --spp 64 --reps 1 > $O/new.log 2>&1
OM_LIB_PATH=raytracingoneweekend_b200/_variants/libom_tiles.so ncu --metrics $M --clock-control none -k regex:k_render --csv --log-file $O/old.csv python tools/prof_run.py C2 --spp 64 --reps 1 > $O/old.log 2>&1
echo done
