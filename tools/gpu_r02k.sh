# round-2 evidence of the current build: parity tests, smoke, both bench arms, other workloads, launch list, full ncu captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_k20.json 2> gpurun_out/bench_k20.err; echo "bench k20 rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1; echo "ref rc=$?"
run() { name=$1; shift; timeout 400 python bench.py --no-cpu --no-ppo --no-e2e "$@" > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$?"; }
run lockstep --phase lockstep --steps 192
run config2 --workload config2
run config3 --workload config3
run loads --control loads
run devgen --exogenous device
run config4 --workload config4
if [ -z "$NOPROF" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 16 --warmup 3 --no-cpu --no-ppo --e2e-steps 2 > gpurun_out/ncu_launch.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 42 -c 1 -o gpurun_out/step_peak -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 22 -c 1 -o gpurun_out/step_quiet -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_full2.log 2>&1; echo "full2 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 140 -c 1 -o gpurun_out/step_spread -f python tools/profile_step.py --phase spread --warm 40 --steps 8 > gpurun_out/ncu_full3.log 2>&1; echo "full3 rc=$?"
fi
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/b*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get("roofline") or {}
        print(f, "%.4g" % d["value"], "ms %.4f" % d["ms_per_step"], "kernel %.4f" % r.get("kernel_ms", 0), "frac %.3f" % r.get("frac", 0),
              "dayfrac %.3f" % r.get("day_mean_frac", 0), "e2e", (d.get("e2e") or {}).get("value"), "packed", (d.get("e2e_packed") or {}).get("value"),
              "ppo", (d.get("ppo_loop") or {}).get("value"), "err", d.get("hub_error_bits"))
    except Exception as e:
        print(f, "unreadable", e)
PY
