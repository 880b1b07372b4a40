# compute-sanitizer over the parity tests that exercise every kernel path at small size: per-warp shared-memory rings,
# __syncwarp hand-offs and shared atomics of the step kernel (heterogeneous classes incl. constant power), the dispatch
# kernel (scalar-load step), auto-reset, the persistent queue.  Summaries go to gpurun_out/sanitize_*.txt.
mkdir -p gpurun_out
T="tests/test_gpu_parity.py::test_heterogeneous_hubs tests/test_gpu_loads.py tests/test_gpu_parity.py::test_masked_reset tests/test_gpu_api.py::test_graph_rollout_equals_stepwise"
for tool in memcheck racecheck initcheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 97 python -m pytest $T -q -x > gpurun_out/sanitize_$tool.txt 2>&1
  echo "$tool rc=$?"; grep -a "ERROR SUMMARY\|RACECHECK SUMMARY\|passed\|failed" gpurun_out/sanitize_$tool.txt | tail -4
done
