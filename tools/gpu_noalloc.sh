# chub_step_loads_static: its test alone in a fresh process, then after all the other GPU tests (pytest -x like the round's end)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "all rc=$?"; tail -4 gpurun_out/gpu_tests.log
