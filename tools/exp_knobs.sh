#!/bin/bash
# sweep the step-kernel launch knobs; prints value / ms_per_step / roofline frac per variant
for b in ${VARIANTS:-12808 12810 12812 25604 25605 25606 51202 51203}; do for y in ${SYNCS:-0}; do
  CHUB_BLOCK=$b CHUB_PHASE_SYNC=$y python bench.py --no-cpu --steps 192 --e2e-steps 2 | tail -1 | python -c "
import sys, json; d=json.loads(sys.stdin.read()); print('variant=$b sync=$y', '%.4g'%d['value'], '%.4f ms'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
done; done
