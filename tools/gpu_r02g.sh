mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_checked_build.py -q -x > gpurun_out/checked.log 2>&1; echo "checked rc=$?"; tail -5 gpurun_out/checked.log
run() { name=$1; shift; timeout 300 env "$@" python bench.py --no-cpu --no-ppo --no-e2e $EXTRA > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$?"; }
EXTRA="--phase lockstep --steps 192" run lock_nolpt CHUB_PHASE_SYNC=1
EXTRA="--phase lockstep --steps 192" run lock_lpt A=1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 42 -c 1 -o gpurun_out/r02g_lock_busy -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_lock.log 2>&1; echo "lock rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 22 -c 1 -o gpurun_out/r02g_lock_quiet -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_lockq.log 2>&1; echo "quiet rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/b_lock_*.json")):
    d = json.loads(open(f).read().strip().splitlines()[-1]); r = d["roofline"]
    print(f, "%.4g" % d["value"], "kernel %.4f" % r["kernel_ms"], "day %.4f" % r["lockstep_day"]["mean_ms"])
PY
