#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/j8_ab.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/j8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j8_pytest.log
tail -5 gpurun_out/j8_pytest.log
timeout 300 python tools/ab_time.py --impl default,v3 --k 20,10,30 --reps 7 >> gpurun_out/j8_ab.txt 2>&1
timeout 300 python tools/ab_time.py --impl default --k 10 --cloud frames --reps 5 >> gpurun_out/j8_ab.txt 2>&1
timeout 300 python tools/quick_time.py 1000000 20 >> gpurun_out/j8_ab.txt 2>&1
cat gpurun_out/j8_ab.txt
