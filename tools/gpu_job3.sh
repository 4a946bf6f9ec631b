#!/bin/bash
# variants of the tile kernel (tools/build_variant.sh), same process layout for each: one ab_time run per library
mkdir -p gpurun_out; : > gpurun_out/j3_ab.txt
for v in "" pred63 nosync pred63nosync $EXTRA_VARIANTS; do
  if [ -z "$v" ]; then lib=pointcloudprocessing_b200/libpcp.so; else lib=tools/_tmp/libpcp_$v.so; fi
  [ -f $lib ] || continue
  echo "== variant ${v:-base}" >> gpurun_out/j3_ab.txt
  PCP_LIB=$PWD/$lib timeout 300 python tools/ab_time.py --impl default --k 20,10,30 --reps 7 >> gpurun_out/j3_ab.txt 2>&1
  PCP_LIB=$PWD/$lib timeout 300 python tools/ab_time.py --impl default --k 10 --cloud frames --reps 5 >> gpurun_out/j3_ab.txt 2>&1
done
cat gpurun_out/j3_ab.txt
