#!/usr/bin/env bash
# Two-GPU half of DESIGN.md §8b.
#   gpurun --gpus 2 --timeout 900 -- 'bash bench/runbook_2gpu.sh'
set -u
out=gpurun_out/runbook2; mkdir -p "$out"
step() {
  local name=$1 t=$2; shift 2
  local t0=$SECONDS
  if timeout "$t" "$@" > "$out/$name.log" 2>&1; then echo "PASS $name ($((SECONDS - t0)) s)"; else echo "FAIL $name (exit $?, $((SECONDS - t0)) s)"; fi | tee -a "$out/summary.txt"
}
: > "$out/summary.txt"
export PYTHONPATH=.
step doctor        120 python -m fluxmpi_b200.doctor --json
step multi_tests   420 env FMPI_TEST_UNVERIFIED=1 python -m pytest tests/test_gpu_multi.py -x -q
step sweep_register 180 python -m fluxmpi_b200.launch -n 2 bench/sweep.py --register --ops allreduce --max-mb 256 --out "$out/sweep_register_2.json"
step sweep_plain    180 python -m fluxmpi_b200.launch -n 2 bench/sweep.py --ops allreduce --max-mb 256 --out "$out/sweep_plain_2.json"
step refshape_2     300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --impl refshape --gpus 2 --steps 5 --warmup 3
step sim_nodes      180 python -m fluxmpi_b200.launch -n 2 --sim-nodes 2 bench/sweep.py --max-mb 64 --ops allreduce --out "$out/sweep_simnodes_2.json"
cat "$out/summary.txt"
