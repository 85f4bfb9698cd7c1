#!/bin/bash
# Multi-GPU evidence in one call: N-GPU bench (ours), all-reduce sweep, toy MLP DP/DDP, two DDP tests.
# usage: bash bench/final_multi.sh N [skip_tests]
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29601 bench.py --gpus $N --steps 15 --warmup 5 2> gpurun_out/scale_ours_$N.err | tail -1 > gpurun_out/scale_ours_$N.json
cut -c1-260 gpurun_out/scale_ours_$N.json
timeout 300 $TR --master-port 29602 bench/allreduce_sweep.py --max-mb 256 > gpurun_out/sweep$N.log 2>&1
tail -12 gpurun_out/sweep$N.log | cut -c1-200
timeout 200 $TR --master-port 29603 bench/toy_bench.py --mode ddp 2>/dev/null | tail -1 | cut -c1-400
timeout 200 python bench/toy_bench.py --mode dp --gpus $N 2>/dev/null | tail -1 | cut -c1-400
if [ -z "$2" ]; then
  timeout 400 python -m pytest tests/test_ddp_gpu.py -x -q -k "gradient_views_with_native or fused_collectives" 2>&1 | tail -3
fi
