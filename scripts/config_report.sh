#!/bin/bash
# One gpurun call: every BASELINE.json config at its full per-GPU size, both math modes, with the unmodified reference timed
# on the box's host cores beside it.  Outputs -> gpurun_out/configs_*.jsonl (copied to profiles/ by hand).
set -u
mkdir -p gpurun_out
timeout 900 python bench_cpu.py c1 c3 c4 c5 > gpurun_out/configs_cpu.jsonl 2> gpurun_out/configs_cpu.err; echo "cpu rc=$?"; cat gpurun_out/configs_cpu.jsonl
for m in fp32 tf32; do
  : > gpurun_out/configs_$m.jsonl
  timeout 300 python scripts/bench_stages.py --stage c1 --mode $m >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 600 python scripts/bench_stages.py --stage c3 --mode $m --batch-scale 1 --iters 3 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c4 --mode $m --batch-scale 1 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c4 --mode $m --batch-scale 0.125 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c5 --mode $m --batch-scale 8 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  echo "== $m rc=$?"; cat gpurun_out/configs_$m.jsonl; tail -2 gpurun_out/configs_$m.err
done
