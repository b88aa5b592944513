#!/bin/bash
# A/B of MUFU.TANH gate activations in the fp16-operand recurrent kernels: error against the CPU reference + timings
set -u
out=gpurun_out/rtc_tanh_err.txt; : > $out
for v in 0 1; do
  touch nntoolkitcore_b200/csrc/recurrent_tc.cu nntoolkitcore_b200/csrc/recurrent_cluster_h16.cu
  NNTK_NVCC_EXTRA="-DRTC_MUFU_TANH=$v -DHT_MUFU_TANH=$v" python -m nntoolkitcore_b200.build > /dev/null || echo "build failed" >> $out
  echo "== MUFU_TANH=$v" >> $out
  timeout 400 python scripts/rtc_tanh_err.py >> $out 2>&1
  for i in 1 2; do
    timeout 120 python scripts/bench_stages.py --stage c4 --mode tf32 --batch-scale 1 --iters 3 2>/dev/null | grep 2xLSTM | cut -c1-130 >> $out
    timeout 120 python scripts/bench_stages.py --stage c5 --mode tf32 --batch-scale 8 --iters 3 2>/dev/null | grep "gru\|chain" | cut -c1-130 >> $out
    timeout 120 python scripts/bench_stages.py --stage c5 --mode tf32 --batch-scale 3.75 --iters 3 2>/dev/null | grep "gru\|chain" | cut -c1-130 >> $out
  done
done
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -5 >> $out
cat $out
