mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_curves.py > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/gpu_tests.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --no-cpu --no-ppo --no-e2e --steps 96 $EXTRA > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$?"; }
EXTRA="--phase lockstep --steps 192" run lock_persist A=1
EXTRA="--phase lockstep --steps 192" run lock_static CHUB_STATIC_GRID=1
EXTRA="--phase spread" run spread28 A=1
EXTRA="--phase spread" run spread28_static CHUB_STATIC_GRID=1
EXTRA="--phase spread --steps 20" run spread28_k20 A=1
EXTRA="--phase spread-quarters" run spread7 A=1
EXTRA="--phase spread --workload config3" run c3_spread A=1
EXTRA="--phase lockstep --workload config3 --steps 192" run c3_lock A=1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/b_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d["roofline"]
        print(f, "%.4g" % d["value"], "ms %.4f" % d["ms_per_step"], "kernel %.4f" % r["kernel_ms"], "frac %.3f" % r["frac"],
              "day %.4f" % r.get("lockstep_day", {}).get("mean_ms", 0), "busy %.4f" % r.get("lockstep_day", {}).get("busy_hours_ms", 0), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
