#!/bin/bash
# One `ncu --set full` capture of the step's kernels (first eager step: every kernel once) for a bench config.
# usage (on the GPU box): tools/ncu_capture.sh <config> <out-dir> [kernel-regex] [max-kernels]
# Only after `python bench.py --config <config> --steps 2 --warmup 1 ...` has exited 0 without ncu.
cfg=$1; out=$2; cnt=${4:-40}; rx=${3:-"tc_lstm|tc_gemm|attn_fused|routing|adam_kernel|ntn_head|embed|smp_|sampler_gather|lstm256|lstm_pack"}
mkdir -p "$out"
args="--config $cfg --steps 2 --warmup 1 --no-cpu-baseline --no-filled --no-graph"
python bench.py $args > "$out/plain_c$cfg.out" 2> "$out/plain_c$cfg.err" || { echo "plain run failed"; exit 1; }
timeout -s KILL 900 ncu --set full --clock-control none --import-source on --kernel-name "regex:$rx" -c $cnt \
  -f -o "$out/full_c$cfg" python bench.py $args > "$out/ncu_c$cfg.log" 2>&1
ls -la "$out"; du -sm "$out"
