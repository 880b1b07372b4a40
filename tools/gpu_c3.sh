mkdir -p gpurun_out; rm -f gpurun_out/c3_*.json
run() { name=$1; shift; timeout 400 env "$@" python bench.py --no-cpu --no-ppo --no-e2e --workload ${WL:-config3} --steps 96 $EXTRA > gpurun_out/c3_$name.json 2> gpurun_out/c3_$name.err; echo "$name rc=$?"; }
for b in 1220 1020 2010 2410; do run size_b$b CHUB_BLOCK=$b; done
WL=config5 run c5_b1420 CHUB_BLOCK=1420
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c3_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); r = d["roofline"]; day = r.get("lockstep_day") or {}
        print(f, "%.4g" % d["value"], "kernel %.4f" % r["kernel_ms"], "frac %.3f" % r["frac"], "day %.4f" % day.get("mean_ms", 0), "busy %.4f" % day.get("busy_hours_ms", 0), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
