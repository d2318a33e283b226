#!/usr/bin/env bash
# One-GPU half of DESIGN.md §8b: first execution of the pieces written without GPU time.
#   gpurun --timeout 900 -- 'bash bench/runbook_1gpu.sh'
# Every step has its own timeout and log under gpurun_out/runbook/; a failing step does not stop
# the later ones (the summary at the end says which passed).
set -u
out=gpurun_out/runbook; mkdir -p "$out"
step() { # name, timeout seconds, command...
  local name=$1 t=$2; shift 2
  local t0=$SECONDS
  if timeout "$t" "$@" > "$out/$name.log" 2>&1; then echo "PASS $name ($((SECONDS - t0)) s)"; else echo "FAIL $name (exit $?, $((SECONDS - t0)) s)"; fi | tee -a "$out/summary.txt"
}
: > "$out/summary.txt"
step doctor            120 python -m fluxmpi_b200.doctor --json
step conv_bn_tests     240 env FMPI_TEST_UNVERIFIED=1 python -m pytest tests/test_gpu_conv_bn.py -x -q
step conv_bn_bench_1sm 240 python bench/conv_bn_bench.py --variant 0 --out "$out/conv_bn_bench_1sm.json"
step conv_bn_bench_2sm 240 python bench/conv_bn_bench.py --variant 1 --out "$out/conv_bn_bench_2sm.json"
step bench_default     180 python bench.py --steps 20 --warmup 5
step bench_conv_bn     180 python bench.py --steps 20 --warmup 5 --conv-bn
step bench_conv_bn_2sm 180 env FMPI_CONV_BN_VARIANT=1 python bench.py --steps 20 --warmup 5 --conv-bn
cat "$out/summary.txt"
