#!/bin/bash
# round-2 GPU job 2: ncu capture of the search kernels (after the same command exited 0 without ncu), full-size tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py -x -q > gpurun_out/j2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j2_pytest.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/j2_plain.json 2> gpurun_out/j2_plain.err || exit 1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"knn_tile5_kernel|search_pca_worklist_kernel" -c 2 -o gpurun_out/r02a_prof -f python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/j2_ncu.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02a_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/j2_ncu2.log 2>&1
tail -3 gpurun_out/j2_pytest.log; ls -la gpurun_out/r02a*
