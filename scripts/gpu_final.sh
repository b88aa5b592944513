#!/bin/bash
# One gpurun call for the round-end evidence: full GPU suite, both bench arms, every config in both math modes,
# launch list of the bench command, full captures of the two recurrent tensor-core kernels.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench_ref.json gpurun_out/bench.json
timeout 600 python bench_cpu.py f1 > gpurun_out/configs_cpu_f1.jsonl 2>/dev/null; cat gpurun_out/configs_cpu_f1.jsonl
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for s in f1 f2 f4; do timeout 300 python scripts/bench_stages.py --stage $s --iters 5; done > gpurun_out/configs_next.jsonl 2>&1; cat gpurun_out/configs_next.jsonl
for m in tf32 x3 fp32; do
  : > gpurun_out/configs_$m.jsonl
  timeout 300 python scripts/bench_stages.py --stage c1 --mode $m >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 600 python scripts/bench_stages.py --stage c3 --mode $m --batch-scale 1 --iters 3 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c4 --mode $m --batch-scale 1 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c4 --mode $m --batch-scale 0.125 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c5 --mode $m --batch-scale 8 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  timeout 900 python scripts/bench_stages.py --stage c5 --mode $m --batch-scale 3.75 --iters 2 >> gpurun_out/configs_$m.jsonl 2>> gpurun_out/configs_$m.err
  echo "== $m rc=$?"; cat gpurun_out/configs_$m.jsonl; tail -2 gpurun_out/configs_$m.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
cap() { # name kernel-regex skip stage mode scale
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/$1 \
    python scripts/bench_stages.py --stage $4 --mode $5 --batch-scale $6 --iters 1 > gpurun_out/$1.log 2>&1; echo "$1 rc=$?"
}
cap lstm_tc recurrent_tc_step 40 c4 tf32 1
cap gru_h16 recurrent_cluster_h16 1 c5 tf32 3.75
cap conv_tc gemm_tf32_kernel 4 c3 tf32 0.25
cap proj_bf16 gemm_tf32_kernel 1 c4 tf32 0.25
cap logmel spectrogram512 9 f1 fp32 0.25
cap dft512 dft512 3 f4 fp32 0.25
ls -la gpurun_out | head -40
