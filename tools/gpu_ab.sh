#!/bin/bash
# A/B of builds of the library (build/ab/<name>.so) on the throughput case, the P/B-slice case and the profile case
# usage: tools/gpu_ab.sh <tag> <name> [<name> ...]
O=gpurun_out; T=$1; shift
cp avrecode_b200/libavrecode_b200.so build/ab/_keep.so
for v in "$@"; do
  cp build/ab/$v.so avrecode_b200/libavrecode_b200.so
  echo "== $v"
  THROUGHPUT_RESIDENT=1 timeout 600 python tools/throughput_case.py ${AB_CLIPS:-32} 2>&1 | tail -1 | tee $O/${T}_tp_$v.txt
  timeout 300 python tools/pb_case.py 2>&1 | tail -1 | tee $O/${T}_pb_$v.txt
  timeout 300 python tools/profile_case.py 30 0 2>&1 | head -1 | tee $O/${T}_profile_$v.txt
done
cp build/ab/_keep.so avrecode_b200/libavrecode_b200.so
