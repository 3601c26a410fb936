#!/bin/bash
# Round-end evidence in one gpurun call: launch list of the bench command, `--set full` capture of the bench's stream
# (application replay: the job hand-out is dynamic, kernel replay gave NaNs), exported as CSV on the box.
# usage: tools/gpu_round2.sh r2z
T=${1:-r2z}; O=gpurun_out
export AVR_BENCH_CORPUS=0 AVR_BENCH_PER_CONFIG=0 AVR_BENCH_FILE_MODE=0
set -x
timeout 400 python bench.py --steps 2 --warmup 3 > $O/${T}_bench_short.json 2> $O/${T}_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 > $O/${T}_ncu_launches.log 2>&1
timeout 300 python tools/profile_case.py 300 32768 > $O/${T}_profile_case.log 2>&1 && \
PROFILE_REPS=1 timeout 1500 ncu --set full --clock-control none --import-source on --replay-mode application -k regex:avr_job_kernel \
    -o /tmp/${T}_full -f python tools/profile_case.py 300 32768 > $O/${T}_ncu_full.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > $O/${T}_full_raw.csv 2> $O/${T}_raw.err
ncu -i /tmp/${T}_full.ncu-rep --page details > $O/${T}_full_details.txt 2>> $O/${T}_raw.err
ncu -i /tmp/${T}_full.ncu-rep --page source --csv 2> $O/${T}_src.err | gzip -9 > $O/${T}_job_source.csv.gz
ls -la $O | tail -12; tail -3 $O/${T}_ncu_full.log
