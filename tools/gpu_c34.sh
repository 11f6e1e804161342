#!/bin/bash
set -u
for c in c2 c3 c4; do python bench.py --config $c --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$c', round(d['ms_per_step'],3),'ms', round(d['value'],1), 'frac', round(d['roofline']['frac'],4), d['config']['workload'][:60])"; done
for v in 20 26 32 40; do AFX_ITEM_STEPS=$v python bench.py --config c3 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('c3 item_steps $v', round(d['ms_per_step'],3),'ms', round(d['value'],1))"; done
