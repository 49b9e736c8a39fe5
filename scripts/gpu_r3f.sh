#!/bin/bash
# r3f (gpurun --gpus 8): the driver's scaling run with the final build: config 2 at N = 1, 2, 4, 8 (ranks aligned
# before the untimed steps), config 3 at N = 8
mkdir -p gpurun_out
O=gpurun_out
: > $O/r3f_scale_config2.jsonl
P=29700
for n in 1 2 4 8; do
  P=$((P+1))
  if [ $n -eq 1 ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline >> $O/r3f_scale_config2.jsonl 2> $O/r3f_n$n.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline >> $O/r3f_scale_config2.jsonl 2> $O/r3f_n$n.err; fi
  echo "n=$n exit $?"
done
P=$((P+1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $P bench.py --config 3 --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > $O/r3f_config3_n8.json 2> $O/r3f_config3_n8.err
python - <<PY
import json
rows=[json.loads(l) for l in open("$O/r3f_scale_config2.jsonl") if l.strip().startswith("{")]
v1=rows[0]["value"]
for d in rows: print("config 2 N=%d: %.2f us  %.3e audio-s/s  eff %.3f  e2e %.3e" % (d["n_gpus"], d["ms_per_step"]*1e3, d["value"], d["value"]/(d["n_gpus"]*v1), d["e2e"]["value"]))
d=json.load(open("$O/r3f_config3_n8.json")); print("config 3 N=8: %.2f us %.3e id %s" % (d["ms_per_step"]*1e3, d["value"], d["mask_identity"]["sha1"]))
PY
