#!/bin/bash
# round-2 GPU job 1: tests, A/B of the tile kernels, bench line, memcheck
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/j1_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/j1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j1_pytest.log
timeout 300 python tools/ab_time.py --impl default,tile4,v3 --k 20,10,30 > gpurun_out/j1_ab.txt 2>&1
timeout 300 python tools/ab_time.py --impl default --k 20 --cells 0.85,0.92,1.0,1.08,1.16,1.25 >> gpurun_out/j1_ab.txt 2>&1
timeout 300 python tools/ab_time.py --impl default,tile4 --k 10 --cloud frames >> gpurun_out/j1_ab.txt 2>&1
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/j1_bench.json 2> gpurun_out/j1_bench.err; echo "bench rc=$?" >> gpurun_out/j1_bench.err
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/memcheck_small.py > gpurun_out/j1_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/j1_memcheck.log
tail -3 gpurun_out/j1_pytest.log; cat gpurun_out/j1_ab.txt; tail -c 600 gpurun_out/j1_bench.json; tail -3 gpurun_out/j1_memcheck.log
