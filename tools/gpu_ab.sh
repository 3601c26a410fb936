#!/bin/bash
# A/B of two builds of the library (build/ab/base.so, build/ab/new.so) on the P/B-slice case and the profile case
O=gpurun_out; T=${1:-ab}
for v in base new; do
  cp build/ab/$v.so avrecode_b200/libavrecode_b200.so
  echo "== $v"
  timeout 300 python tools/pb_case.py 2>&1 | tail -4 | tee $O/${T}_pb_$v.txt
  timeout 300 python tools/profile_case.py 30 0 2>&1 | tail -3 | tee $O/${T}_profile_$v.txt
done
cp build/ab/new.so avrecode_b200/libavrecode_b200.so
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/${T}_pytest.txt
