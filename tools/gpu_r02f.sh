mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/gpu_tests.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --no-cpu --no-ppo --no-e2e $EXTRA > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$?"; }
EXTRA="--steps 96" run spread A=1
EXTRA="--phase lockstep --steps 192" run lock A=1
EXTRA="--phase lockstep --steps 192" run lock_static CHUB_STATIC_GRID=1
EXTRA="--workload config3 --steps 96" run c3_sorted A=1
EXTRA="--workload config3 --steps 96" run c3_unsorted CHUB_BENCH_UNSORTED=1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/b_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d["roofline"]
        print(f, "%.4g" % d["value"], "ms %.4f" % d["ms_per_step"], "kernel %.4f" % r["kernel_ms"], "frac %.3f" % r["frac"],
              "day %.4f" % r.get("lockstep_day", {}).get("mean_ms", 0), "busy %.4f" % r.get("lockstep_day", {}).get("busy_hours_ms", 0), "dayfrac %.3f" % r.get("day_mean_frac", 0), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
