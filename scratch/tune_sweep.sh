#!/bin/bash
for r in 1 2; do
  KGE_SWEEP_RING=$r timeout 120 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --train-only 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); s=d['stages_ms_per_step']; print('ring=$r step=%.4f entity=%.4f sweep=%.4f' % (d['ms_per_step'], s['entity_gather+adam'], s['fused_sweep(scores+loss+dL/dq)']))"
done
