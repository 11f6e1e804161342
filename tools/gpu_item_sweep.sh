#!/bin/bash
# item-size sweep of the streaming kernel (AFX_ITEM_STEPS; default = chosen per batch, at most 32)
set -u
for n in 10000 2000; do
for v in 32 48 64 96 127; do
  if [ $v -eq 0 ]; then unset AFX_ITEM_STEPS; else export AFX_ITEM_STEPS=$v; fi
  python bench.py --clips $n --steps 6 --warmup 3 --no-e2e --no-cpu-baseline --no-f0 --no-legs 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('clips $n item_steps $v: ms', round(d['ms_per_step'],3), 'value', round(d['value'],1))"
done; done
