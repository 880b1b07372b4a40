# parity soaks of the round-2 build (new seeds): default hubs, heterogeneous + scalar-load steps, host path, device generator
mkdir -p gpurun_out; out=gpurun_out/parity_soak.txt
echo "# tools/parity_soak.py on B200 (round r02k build: constant-bank table, single-body stream walk, self-advancing step index): GPU engine vs CPU checker, all integer fields + fp32 pile state" > $out
run() { echo "## python tools/parity_soak.py $*" >> $out; timeout 900 python tools/parity_soak.py "$@" 2>&1 | tail -1 >> $out; }
run --hubs 65536 --days 20 --seed 1201 --every 16
run --hubs 65536 --days 10 --classes 128 --seed 1202 --every 16 --loads-every 7
run --hubs 20000 --days 6 --host --classes 16 --seed 1203
run --hubs 16384 --days 10 --exogenous device --seed 1204
run --hubs 131072 --days 5 --seed 1205 --every 16
cat $out
