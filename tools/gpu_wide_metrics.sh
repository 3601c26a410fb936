#!/bin/bash
# The wide decompress kernel's counters come back NaN from ncu's kernel replay (its blocks spin on rings other warps fill);
# application replay re-runs the whole case per pass instead.  usage: tools/gpu_wide_metrics.sh r1k
T=${1:-rX}; O=gpurun_out
M=dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum
M=$M,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio
M=$M,smsp__average_warp_latency_per_inst_issued.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
M=$M,launch__registers_per_thread,launch__shared_mem_per_block_dynamic,sm__warps_active.avg.pct_of_peak_sustained_active,sm__throughput.avg.pct_of_peak_sustained_elapsed
set -x
PROFILE_REPS=1 timeout 500 ncu --replay-mode application --metrics $M --clock-control none --csv --page raw \
    --log-file $O/${T}_app_replay_raw.csv python tools/profile_case.py 300 32768 > $O/${T}_app_replay.log 2>&1
tail -3 $O/${T}_app_replay.log; ls -la $O
