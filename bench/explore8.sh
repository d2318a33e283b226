o=gpurun_out/c11; mkdir -p $o
L="python -m fluxmpi_b200.launch -n 8"
FMPI_PIPE_TRACE=1 timeout 120 $L bench/pipe_probe.py --mb 256 --op allreduce > $o/probe_ar.log 2>&1; head -12 $o/probe_ar.log; grep -A6 "role link" $o/probe_ar.log; grep -A6 "role out" $o/probe_ar.log
FMPI_PIPE_TRACE=1 timeout 120 $L bench/pipe_probe.py --mb 256 --op bcast > $o/probe_bc.log 2>&1; head -8 $o/probe_bc.log; grep -A5 "role link" $o/probe_bc.log; grep -A5 "role out" $o/probe_bc.log
FMPI_PIPE_TRACE=1 timeout 120 $L bench/pipe_probe.py --mb 256 --op reduce > $o/probe_rd.log 2>&1; head -8 $o/probe_rd.log; grep -A5 "role link" $o/probe_rd.log; grep -A5 "role out" $o/probe_rd.log
for cfg in "FMPI_PIPE_CHUNK_KB=8192" "FMPI_PIPE_LINK_NVLS=8 FMPI_PIPE_COPY_NVLS=96" "FMPI_PIPE_LINK_NVLS=24 FMPI_PIPE_COPY_NVLS=112" "FMPI_PIPE_SLOTS=4 FMPI_PIPE_CHUNK_KB=12288"; do
  for op in allreduce bcast reduce; do env $cfg timeout 120 $L bench/pipe_probe.py --mb 256 --op $op 2>&1 | grep busbw; done; done
timeout 300 $L bench/fused_bench.py --iters 5 --models resnet50,gpt2_medium --blocks 48,96,128,160 --out $o/fused_8.json > $o/fused8.log 2>&1; grep "^{" $o/fused8.log
timeout 400 $L bench/sweep.py --out $o/sweep_8.json > $o/sweep8.log 2>&1; echo "sweep rc=$?"
python - <<PY
import json
for l in open("gpurun_out/c11/sweep8.log"):
    if l.startswith("{"):
        r=json.loads(l); print(r["bytes"], {k.replace("_busbw_gbs","").replace("allreduce","ar").replace("bcast","bc").replace("reduce","rd"):v for k,v in r.items() if k.endswith("busbw_gbs") or (k.endswith("_us") and r["bytes"]<=65536)})
PY
