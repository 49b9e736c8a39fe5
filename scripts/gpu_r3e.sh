#!/bin/bash
# r3e (gpurun --gpus 2): does aligning the ranks before the untimed steps remove the N > 1 loss?
mkdir -p gpurun_out
O=gpurun_out
P=29600
for i in 1 2 3; do
  for mode in align noalign; do
    P=$((P+1))
    if [ $mode = noalign ]; then export B200MEL_BENCH_NO_ALIGN=1; else unset B200MEL_BENCH_NO_ALIGN; fi
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3e_${mode}_$i.json 2> $O/r3e_${mode}_$i.err
    python - <<PY
import json
try:
    d=json.load(open("$O/r3e_${mode}_$i.json")); print("N=2 $mode #$i: %.2f us  %.3e" % (d["ms_per_step"]*1e3, d["value"]))
except Exception as e: print("$mode failed", e)
PY
  done
done
unset B200MEL_BENCH_NO_ALIGN
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $O/r3e_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('$O/r3e_n1.json')); print('N=1: %.2f us' % (d['ms_per_step']*1e3))"
