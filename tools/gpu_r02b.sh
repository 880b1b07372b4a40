# why is a de-phased batch slow?  full ncu captures of one launch: per-warp spread, per-block spread, lock-step busy hour
mkdir -p gpurun_out
for ph in spread spread-blocks; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 100 -c 1 -o gpurun_out/r02_$ph -f python tools/profile_step.py --phase $ph --warm 4 --steps 2 > gpurun_out/ncu_$ph.log 2>&1; echo "$ph rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s 42 -c 1 -o gpurun_out/r02_lockstep_busy -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_lock.log 2>&1; echo "lock rc=$?"
ls -la gpurun_out/*.ncu-rep
