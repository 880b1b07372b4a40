mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default rc=$?"; tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --workload config4 --no-cpu > gpurun_out/bench_config4.json 2> gpurun_out/bench_config4.err; echo "config4 rc=$?"; tail -3 gpurun_out/bench_config4.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/bench_ref.err
timeout 300 python bench.py --workload config2 --no-cpu --no-e2e > gpurun_out/bench_config2.json 2> gpurun_out/bench_config2.err; echo "config2 rc=$?"; tail -3 gpurun_out/bench_config2.err
timeout 300 python bench.py --control loads --no-cpu --no-e2e > gpurun_out/bench_loads.json 2> gpurun_out/bench_loads.err; echo "loads rc=$?"; tail -3 gpurun_out/bench_loads.err
timeout 300 python bench.py --exogenous device --no-cpu --no-e2e --no-ppo > gpurun_out/bench_devgen.json 2> gpurun_out/bench_devgen.err; echo "devgen rc=$?"; tail -3 gpurun_out/bench_devgen.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
