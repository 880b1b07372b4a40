mkdir -p gpurun_out
for h in 131072 262144 524288 1048576 2097152; do
python bench.py --no-cpu --no-ppo --no-e2e --phase lockstep --steps 192 --hubs $h 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; day=r['lockstep_day']; c=day['curve_ms']
print('hubs $h: ms/step %.4f kernel_ms %.4f day %.4f min %.4f busy %.4f  ns/hub %.4f' % (d['ms_per_step'], r['kernel_ms'], day['mean_ms'], min(c), day['busy_hours_ms'], 1e6*day['mean_ms']/$h))"
done
