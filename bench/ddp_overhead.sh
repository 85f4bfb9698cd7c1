#!/bin/bash
# Where does the multi-GPU step lose time against the single-GPU step?  Runs bench.py at N GPUs with
# a few communication settings; results land in gpurun_out/ddp_overhead_N.txt.
N=${1:-2}
OUT=gpurun_out/ddp_overhead_$N.txt
: > $OUT
run() {
  echo "== $*" >> $OUT
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 15 --warmup 5 "$@" 2>/dev/null | tail -1 | python -c '
import json,sys
d=json.loads(sys.stdin.read()); print(d["ms_per_step"], "ms/step", d["value"], "img/s  e2e", d["e2e"]["ms_per_step"], "ms", d["info"]["ddp"]["bucket_modes"])' >> $OUT 2>&1
}
run
run --allreduce nccl
run --no-broadcast-buffers
run --comm-blocks 8
run --comm-blocks 64
run --comm-dtype bf16
run --allreduce oneshot
cat $OUT
