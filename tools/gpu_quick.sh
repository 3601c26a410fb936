#!/bin/bash
# quick GPU check: tests + floors.  usage: tools/gpu_quick.sh <tag> [pytest -k expr]
T=${1:-rq}; O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q ${2:+-k "$2"} > $O/${T}_pytest.txt 2>&1; tail -15 $O/${T}_pytest.txt
timeout 120 python tools/probe_case.py 20000 > $O/${T}_probes.txt 2>&1; cat $O/${T}_probes.txt
timeout 300 python tools/profile_case.py 30 0 > $O/${T}_profile_case.log 2>&1; cat $O/${T}_profile_case.log
for K in 2 4 8; do
  PROFILE_MAX_STREAMS=$K timeout 300 python tools/profile_case.py 30 32768 > $O/${T}_profile_case_wide$K.log 2>&1; cat $O/${T}_profile_case_wide$K.log
done
