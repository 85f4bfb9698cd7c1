#!/bin/bash
# Scaling run on an N-GPU box: ours and the reference arm at the given GPU counts, all-reduce sweep,
# multi-GPU tests.  usage: bench/run_scaling.sh "8 4" [steps]
COUNTS=${1:-"8 4 2 1"}
STEPS=${2:-20}
mkdir -p gpurun_out
port=29600
for n in $COUNTS; do
  for impl in reference ours; do
    port=$((port+1))
    if [ "$n" = "1" ]; then
      timeout 300 python bench.py --impl $impl --gpus 1 --steps $STEPS --warmup 5 > gpurun_out/scale_${impl}_${n}.json 2> gpurun_out/scale_${impl}_${n}.err
    else
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --impl $impl --gpus $n --steps $STEPS --warmup 5 > gpurun_out/scale_${impl}_${n}.json 2> gpurun_out/scale_${impl}_${n}.err
    fi
    echo "$impl N=$n rc=$? $(cut -c1-230 gpurun_out/scale_${impl}_${n}.json)"
  done
done
