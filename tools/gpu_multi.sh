#!/bin/bash
# multi-GPU job: configs[3] strong scaling (tools/cfg3_scaling.py) and the bench line at N GPUs.   usage: gpu_multi.sh <N> <tag>
N=$1; TAG=${2:-r02}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = 1 ]; then RUNC="python"; else RUNC="$RUN --master-port 29511"; fi
timeout 1500 $RUNC tools/cfg3_scaling.py --steps 5 --out gpurun_out/${TAG}_cfg3_${N}gpu.json > /dev/null 2> gpurun_out/${TAG}_cfg3_${N}gpu.err; echo "cfg3 rc=$?"
grep -E "bitwise|points_per_s_wall|ms_per_step" gpurun_out/${TAG}_cfg3_${N}gpu.json; tail -3 gpurun_out/${TAG}_cfg3_${N}gpu.err
if [ "$N" != 1 ]; then
  timeout 900 $RUN --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
  timeout 600 $RUN --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_${N}gpu.json 2>> gpurun_out/${TAG}_bench_${N}gpu.err; echo "ref rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/${TAG}_bench_${N}gpu.json'));print({k:d[k] for k in ('value','ms_per_step','parity_checked','n_gpus')}, d['e2e']['value']); r=json.load(open('gpurun_out/${TAG}_bench_reference_${N}gpu.json')); print('ref cores', r['cpu_baseline']['cores'], r['value'])"
  tail -3 gpurun_out/${TAG}_bench_${N}gpu.err
fi
