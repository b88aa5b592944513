#!/bin/bash
# One gpurun call: ncu --set full captures of the GEMM-shaped and recurrent kernels (one launch each), text details exported.
set -u
mkdir -p gpurun_out
python scripts/measure_peaks.py > gpurun_out/peaks_extra.json 2> gpurun_out/peaks_extra.err; cat gpurun_out/peaks_extra.json
cap() { # name kernel-regex skip stage mode scale
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/$1 \
    python scripts/bench_stages.py --stage $4 --mode $5 --batch-scale $6 --iters 1 > gpurun_out/$1.log 2>&1; echo "$1 rc=$?"
  ncu -i gpurun_out/$1.ncu-rep --page details > gpurun_out/$1.details.txt 2>&1
}
cap conv_tc gemm_tf32_kernel 4 c3 tf32 0.25
cap lstm_tc recurrent_tc_step 40 c4 tf32 0.25
cap gru_ctc recurrent_cluster_tc 1 c5 tf32 1.875
cap gru_cluster recurrent_cluster_kernel 1 c5 fp32 0.25
cap conv_fp32 gemm_fp32_kernel 4 c3 fp32 0.25
cap lstm_fp32 recurrent_step_kernel 40 c4 fp32 0.25
ls -la gpurun_out
