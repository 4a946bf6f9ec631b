#!/bin/bash
# the round's standard single-GPU job: every GPU test, the bench line (after it exited 0 without ncu: launch list and a
# full capture of the search kernels), the parity statistics.   usage: gpu_full.sh <tag>
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/${TAG}_ncu1.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"knn_tile6_kernel|search_pca_worklist_kernel" -c 2 -o gpurun_out/${TAG}_prof -f python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/${TAG}_ncu2.log 2>&1
timeout 900 python tools/parity_stats.py gpurun_out/${TAG}_parity_stats.json > gpurun_out/${TAG}_parity_stats.log 2>&1; tail -2 gpurun_out/${TAG}_parity_stats.log
python -c "import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));print({k:d[k] for k in ('value','ms_per_step','parity_checked')}, d['e2e']['value'], d['roofline']['kernel_ms'])"
