mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/gpu_tests.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --no-cpu --no-ppo --no-e2e $EXTRA > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$?"; }
EXTRA="--phase lockstep --steps 192" run la_lock A=1
EXTRA="--steps 96" run la_spread A=1
EXTRA="--phase lockstep --steps 192" run la_static CHUB_STATIC_GRID=1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/b_la_*.json")):
    d = json.loads(open(f).read().strip().splitlines()[-1]); r = d["roofline"]; c = r["lockstep_day"]["curve_ms"]
    print(f, "%.4g" % d["value"], "kernel %.4f" % r["kernel_ms"], "day %.4f" % r["lockstep_day"]["mean_ms"], "min %.4f" % min(c), "busy %.4f" % r["lockstep_day"]["busy_hours_ms"], "err", d["hub_error_bits"])
PY
