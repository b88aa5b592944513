#!/bin/bash
# Rebuilds the per-step recurrent kernel with different epilogue chunk / prefetch-depth settings and times C4/C5 (TF32).
set -u
mkdir -p gpurun_out
out=gpurun_out/rtc_variants.txt; : > $out
for v in "8 1" "4 3" "4 2" "4 1"; do
  set -- $v
  touch nntoolkitcore_b200/csrc/recurrent_tc.cu
  NNTK_NVCC_EXTRA="-DRTC_CH=$1 -DRTC_PD=$2" python -m nntoolkitcore_b200.build > /dev/null || { echo "build failed $v" >> $out; continue; }
  echo "== CH=$1 PD=$2" >> $out
  timeout 600 python scripts/bench_stages.py --stage c4 --mode tf32 --batch-scale 1 --iters 3 >> $out 2>/dev/null
  timeout 600 python scripts/bench_stages.py --stage c4 --mode tf32 --batch-scale 0.125 --iters 3 >> $out 2>/dev/null
  timeout 600 python scripts/bench_stages.py --stage c5 --mode tf32 --batch-scale 8 --iters 3 >> $out 2>/dev/null
done
touch nntoolkitcore_b200/csrc/recurrent_tc.cu; python -m nntoolkitcore_b200.build > /dev/null
cut -c1-400 $out
