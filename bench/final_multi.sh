#!/bin/bash
# Multi-GPU evidence in one call: the multi-GPU tests, the all-reduce sweep (1 KB - 1 GiB), toy MLP DP / DDP against
# torch, and the N-GPU bench.  usage: bash bench/final_multi.sh N [max_mb]
N=${1:-8}
MAXMB=${2:-1024}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "== pytest (multi-GPU files) on $N GPUs"
timeout 900 python -m pytest tests/test_ddp_gpu.py tests/test_dp_mp_gpu.py -q -p no:cacheprovider -rs > gpurun_out/pytest_multi_$N.log 2>&1
echo "pytest_rc=$?"; tail -8 gpurun_out/pytest_multi_$N.log
echo "== all-reduce sweep"
timeout 400 $TR --master-port 29602 bench/allreduce_sweep.py --max-mb $MAXMB > gpurun_out/allreduce_sweep_$N.md 2> gpurun_out/allreduce_sweep_$N.err
echo "sweep_rc=$?"; grep -v "^\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/allreduce_sweep_$N.md | cut -c1-260
echo "== toy MLP"
timeout 200 $TR --master-port 29603 bench/toy_bench.py --mode ddp 2>/dev/null | tail -1 | cut -c1-500
timeout 200 python bench/toy_bench.py --mode dp --gpus $N 2>/dev/null | tail -1 | cut -c1-500
echo "== bench.py"
timeout 300 $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/scale_ours_$N.err | tail -1 > gpurun_out/scale_ours_$N.json
cut -c1-400 gpurun_out/scale_ours_$N.json
python - <<PY
import json
d = json.load(open("gpurun_out/scale_ours_$N.json"))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, d["e2e"], d["info"]["e2e_sync"], d["info"]["ddp"]["bucket_modes"], d["clocks"])
PY
