# round-2 GPU check: parity suite (all failures shown), curve sweep, bench in both phase modes and launch shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_curves.py > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/gpu_tests.log
timeout 300 python -m pytest tests/test_gpu_curves.py -q -s > gpurun_out/curves.log 2>&1; echo "curves rc=$?"; grep -a "differ\|passed\|failed" gpurun_out/curves.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
for v in 2810 1620 1220 830 2010; do
  CHUB_BLOCK=$v timeout 300 python bench.py --no-cpu --no-ppo --no-e2e --steps 96 > gpurun_out/bench_spread_$v.json 2> gpurun_out/bench_spread_$v.err; echo "spread $v rc=$?"
done
timeout 300 python bench.py --no-cpu --no-ppo --no-e2e --phase spread-blocks --steps 96 > gpurun_out/bench_spread_blocks.json 2>&1; echo "blocks rc=$?"
timeout 300 python bench.py --no-cpu --no-ppo --no-e2e --phase lockstep --steps 192 > gpurun_out/bench_lockstep.json 2>&1; echo "lockstep rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d["roofline"]
        print(f, "%.4g" % d["value"], "ms %.4f" % d["ms_per_step"], "kernel %.4f" % r["kernel_ms"], "frac %.3f" % r["frac"],
              "day %.4f" % r.get("lockstep_day", {}).get("mean_ms", 0), "e2e", (d.get("e2e") or {}).get("value"), "packed", (d.get("e2e_packed") or {}).get("value"),
              "ppo", (d.get("ppo_loop") or {}).get("value"), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
