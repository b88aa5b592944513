#!/bin/bash
# One gpurun call: GPU tests, bench (both arms), per-stage timings.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
for m in fp32 tf32; do
  timeout 900 python scripts/bench_stages.py --stage all --mode $m --batch-scale 0.25 --iters 2 > gpurun_out/stages_$m.jsonl 2> gpurun_out/stages_$m.err; echo "stages $m rc=$?"
  cat gpurun_out/stages_$m.jsonl; tail -3 gpurun_out/stages_$m.err
done
