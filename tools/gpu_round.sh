# one GPU evidence round: parity tests, smoke, both bench arms, the other configs, ncu launch list, full ncu captures (busy + quiet hour)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1; echo "ref rc=$?"
timeout 300 python bench.py --loop ppo --steps 384 --warmup 128 --no-cpu > gpurun_out/bench_ppo.json 2> gpurun_out/bench_ppo.err; echo "ppo rc=$?"
timeout 300 python bench.py --host-threads 0 --no-cpu --steps 48 > gpurun_out/bench_e2e_unpacked.json 2>&1; echo "unpacked rc=$?"
timeout 300 python bench.py --workload config2 --no-cpu > gpurun_out/bench_config2.json 2>&1; echo "config2 rc=$?"
timeout 300 python bench.py --workload config3 --no-cpu > gpurun_out/bench_config3.json 2>&1; echo "config3 rc=$?"
timeout 300 python bench.py --control loads --no-cpu > gpurun_out/bench_loads.json 2>&1; echo "loads rc=$?"
timeout 300 python bench.py --exogenous device --no-cpu > gpurun_out/bench_devgen.json 2>&1; echo "devgen rc=$?"
timeout 100 python tools/single_env_rate.py > gpurun_out/single_env.txt 2>&1; echo "single rc=$?"
if [ -z "$NOPROF" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 16 --warmup 3 --no-cpu --e2e-steps 2 > gpurun_out/ncu_launch.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 42 -c 1 -o gpurun_out/step_peak -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 22 -c 1 -o gpurun_out/step_quiet -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_full2.log 2>&1; echo "full2 rc=$?"
python tools/step_times.py > gpurun_out/step_times.txt 2>&1
fi
