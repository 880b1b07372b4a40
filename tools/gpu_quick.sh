# quick GPU check: parity tests, then a short bench per launch variant (VARIANTS) and phase-sync setting (SYNCS)
mkdir -p gpurun_out
if [ -z "$NOTEST" ]; then timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/gpu_tests.log; fi
for b in ${VARIANTS:-0}; do for y in ${SYNCS:-0}; do
  CHUB_BLOCK=$b CHUB_PHASE_SYNC=$y timeout 300 python bench.py --no-cpu --steps 192 --e2e-steps 2 2>&1 | tail -1 | python -c "
import sys, json
try:
    d=json.loads(sys.stdin.read()); print('variant=$b sync=$y', '%.4g'%d['value'], '%.4f ms'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'], 'err', d['hub_error_bits'])
except Exception as e: print('variant=$b sync=$y failed', e)"
done; done
