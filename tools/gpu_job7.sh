#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/j7_ab.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/j7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j7_pytest.log
tail -5 gpurun_out/j7_pytest.log
for v in "" ch256 ch64; do
  if [ -z "$v" ]; then lib=pointcloudprocessing_b200/libpcp.so; else lib=tools/_tmp/libpcp_$v.so; fi
  [ -f $lib ] || continue
  echo "== variant ${v:-base}" >> gpurun_out/j7_ab.txt
  PCP_LIB=$PWD/$lib timeout 300 python tools/ab_time.py --impl default --k 20,10,30 --reps 7 >> gpurun_out/j7_ab.txt 2>&1
done
PCP_LIB=$PWD/pointcloudprocessing_b200/libpcp.so timeout 300 python tools/ab_time.py --impl default,tile6,tile4 --k 10 --cloud frames --reps 5 >> gpurun_out/j7_ab.txt 2>&1
cat gpurun_out/j7_ab.txt
