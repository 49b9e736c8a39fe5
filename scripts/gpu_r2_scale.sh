#!/bin/bash
# r2 multi-GPU round (gpurun --gpus 8): configs 3 and 5 at N = 1, 2, 4, 8 (mask identity, real-audio accounting),
# config 2 at N = 8, and the host-link probe at N = 1, 2, 4, 8 (+ core binding / write-combined at 8)
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi topo -m > $O/r2_topo.txt 2>&1
lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)" >> $O/r2_topo.txt 2>&1
PORT=29500
tr() { # n args...
  n=$1; shift
  PORT=$((PORT+1))
  if [ $n -eq 1 ]; then timeout 600 python "$@"; else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT "$@"; fi
}
for c in 3 5; do
  : > $O/r2_scale_config$c.jsonl
  for n in 1 2 4 8; do
    extra="--no-cpu-baseline"; [ $n -eq 1 ] && extra=""
    tr $n bench.py --config $c --gpus $n --steps 20 --warmup 3 $extra >> $O/r2_scale_config$c.jsonl 2> $O/r2_scale_config${c}_n$n.err
    echo "config $c n=$n exit $?"
  done
  python - <<PY
import json
for l in open("$O/r2_scale_config$c.jsonl"):
    try:
        d=json.loads(l); print("config $c N=%d: %.1f us %.3e audio-s/s frac %.4f e2e %s id %s" % (d["n_gpus"], d["ms_per_step"]*1e3, d["value"], d["roofline"]["frac"], (d.get("e2e") or {}).get("value"), (d.get("mask_identity") or {}).get("sha1")))
    except Exception as e: print("bad line", e, l[:100])
PY
done
tr 8 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > $O/r2_bench_8gpu.json 2> $O/r2_bench_8gpu.err; echo "config 2 n=8 exit $?"; cut -c1-400 $O/r2_bench_8gpu.json
: > $O/r2_pcie_concurrent.jsonl
for n in 1 2 4 8; do tr $n experiments/pcie_concurrent.py >> $O/r2_pcie_concurrent.jsonl 2>> $O/r2_pcie.err; done
tr 8 experiments/pcie_concurrent.py --bind >> $O/r2_pcie_concurrent.jsonl 2>> $O/r2_pcie.err
tr 8 experiments/pcie_concurrent.py --wc >> $O/r2_pcie_concurrent.jsonl 2>> $O/r2_pcie.err
tr 8 experiments/pcie_concurrent.py --chunks 8 >> $O/r2_pcie_concurrent.jsonl 2>> $O/r2_pcie.err
tr 4 experiments/pcie_concurrent.py --bind >> $O/r2_pcie_concurrent.jsonl 2>> $O/r2_pcie.err
cat $O/r2_pcie_concurrent.jsonl; tail -5 $O/r2_pcie.err
