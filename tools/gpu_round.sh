#!/bin/bash
# One gpurun call per round end: bench (both arms), launch list, `--set full` capture exported as CSV on the box
# (the .ncu-rep itself is > 64 MiB with --import-source), other BASELINE configs.  usage: tools/gpu_round.sh r1k
T=${1:-rX}; O=gpurun_out
set -x
timeout 400 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
timeout 400 python bench.py --impl reference > $O/${T}_bench_reference.json 2> $O/${T}_bench_reference.err
timeout 300 python tools/profile_case.py 300 32768 > $O/${T}_profile_case.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python tools/profile_case.py 300 32768 > $O/${T}_ncu_launches.log 2>&1
PROFILE_REPS=1 timeout 500 ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
    python tools/profile_case.py 300 32768 > $O/${T}_ncu_full.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > $O/${T}_full_raw.csv 2> $O/${T}_raw.err
ncu -i /tmp/${T}_full.ncu-rep --page source --csv --kernel-name regex:avr_wide_kernel 2> $O/${T}_src.err | gzip -9 > $O/${T}_wide_source.csv.gz
for w in cfg1 cfg3 cfg4; do timeout 300 python bench.py --workload $w > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err; done
ls -la $O; tail -5 $O/${T}_bench.err
