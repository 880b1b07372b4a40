# multi-GPU evidence on one box: weak scaling of the default line (config 5 shard per GPU), strong scaling of config 4
mkdir -p gpurun_out
tr() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n "$@"; }
for n in ${NS:-2 4 8}; do
  tr $n --no-cpu > gpurun_out/scale_config5_${n}gpus.json 2> gpurun_out/scale_config5_${n}gpus.err; echo "config5 x$n rc=$?"
  tr $n --no-cpu --workload config4 > gpurun_out/scale_config4_${n}gpus.json 2> gpurun_out/scale_config4_${n}gpus.err; echo "config4 x$n rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/scale_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); r = d["roofline"]
        print(f, "n", d["n_gpus"], "%.4g" % d["value"], "ms %.4f" % d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), "packed", (d.get("e2e_packed") or {}).get("value"),
              "ppo", (d.get("ppo_loop") or {}).get("value"), "year", (d.get("year") or {}).get("store_within_bounds"), "err", d["hub_error_bits"])
    except Exception as e:
        print(f, "unreadable", e)
PY
