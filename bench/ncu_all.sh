#!/bin/bash
# One `ncu --set full` capture per kernel family (bench/ncu_kernels.py), each preceded by the plain run of the
# same command.  Run under gpurun on ONE GPU; the .ncu-rep files land in gpurun_out/ncu/.
set -u
mkdir -p gpurun_out/ncu
: > gpurun_out/ncu/plain.jsonl
FAMS=${1:-$(python bench/ncu_kernels.py list)}
for f in $FAMS; do
  python bench/ncu_kernels.py $f >> gpurun_out/ncu/plain.jsonl 2> gpurun_out/ncu/$f.err || { echo "plain run of $f failed"; tail -3 gpurun_out/ncu/$f.err; continue; }
  # the script brackets ONE call of the family with cudaProfilerStart/Stop after its warm-up and timing
  timeout 300 ncu --set full --clock-control none --import-source on --profile-from-start off \
      -c 6 -f -o gpurun_out/ncu/$f python bench/ncu_kernels.py $f > gpurun_out/ncu/$f.log 2>&1
  echo "$f ncu_rc=$?"
done
cat gpurun_out/ncu/plain.jsonl
ls -la gpurun_out/ncu/*.ncu-rep | awk '{print $5, $9}'
