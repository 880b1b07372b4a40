# experiment round: parity tests (optional), lock-step day + de-phased bench of the current build
mkdir -p gpurun_out
if [ -z "$NOTEST" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/gpu_tests.log; fi
run() { name=$1; shift; timeout 400 python bench.py --no-cpu --no-ppo --no-e2e "$@" > gpurun_out/x_$name.json 2> gpurun_out/x_$name.err; echo "$name rc=$?"; }
run lock --phase lockstep --steps 192
run spread --steps 96
for w in $EXTRA_WORKLOADS; do run $w --workload $w --steps 96; done
if [ -n "$LOADS" ]; then run loads --control loads --steps 96; fi
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/x_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); r = d["roofline"]; day = r.get("lockstep_day") or {}
        c = day.get("curve_ms") or [0]
        print(f, "%.4g" % d["value"], "kernel %.4f" % r["kernel_ms"], "frac %.3f" % r["frac"], "day %.4f" % day.get("mean_ms", 0),
              "min %.4f" % min(c), "busy %.4f" % day.get("busy_hours_ms", 0), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
