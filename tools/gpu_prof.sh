# one full ncu capture of the step kernel at a busy hour (step 42: all piles occupied) and a quiet one (step 24)
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:chub_step_kernel -s ${SKIP:-42} -c ${COUNT:-1} -o gpurun_out/${NAME:-step_full} -f python tools/profile_step.py --warm 40 --steps 4 > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
