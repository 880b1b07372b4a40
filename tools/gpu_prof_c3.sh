mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 42 -c 1 -o gpurun_out/c3_peak -f python tools/profile_step.py --workload config3 --hubs 262144 --warm 40 --steps 4 > gpurun_out/ncu_c3.log 2>&1; echo "c3 rc=$?"
