for c in 4 8 16 32; do B200MEL_HOST_CHUNKS=$c python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 20 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); print('chunks $c e2e %.3e  kernel %.3e' % (d['e2e']['value'], d['value']))"; done
