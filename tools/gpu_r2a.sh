#!/bin/bash
# r2a: floors (avr_probe) + source-correlated ncu profile of the ORDINARY job kernels on a 30-frame cfg2-shaped case
T=${1:-r2a}; O=gpurun_out
set -x
timeout 120 python tools/probe_case.py 20000 > $O/${T}_probes.txt 2>&1
timeout 300 python tools/profile_case.py 30 0 > $O/${T}_profile_case.log 2>&1
PROFILE_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -o /tmp/${T}_full -f \
    python tools/profile_case.py 30 0 > $O/${T}_ncu_full.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > $O/${T}_full_raw.csv 2> $O/${T}_raw.err
ncu -i /tmp/${T}_full.ncu-rep --page source --print-source cuda,sass --csv 2> $O/${T}_src.err | gzip -9 > $O/${T}_source.csv.gz
ls -la $O; cat $O/${T}_probes.txt; cat $O/${T}_profile_case.log
