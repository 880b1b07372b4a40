"""Copy one GPU evidence round from gpurun_out/ (scratch) into profiles/ (tracked) under a prefix, with the ncu captures
reduced to text summaries and the DRAM bytes of the captured launches written where bench.py's roofline.traffic reads them.
Usage: python tools/collect_profiles.py r02j"""
import csv, io, json, os, shutil, subprocess, sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst, pre = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles"), sys.argv[1]


def put(name, out):
    p = os.path.join(src, name)
    if os.path.exists(p) and os.path.getsize(p):
        shutil.copyfile(p, os.path.join(dst, "%s_%s" % (pre, out)))
        return True
    return False


put("gpu_tests.log", "gpu_tests.log")
put("bench.json", "bench_config5.json")
put("bench_k20.json", "bench_config5_driver_window_20_steps.json")
put("bench_ref.json", "bench_reference_arm.json")
put("launches.csv", "launches.csv")
for n, o in (("lockstep", "config5_lockstep"), ("config2", "config2"), ("config3", "config3"), ("loads", "config5_load_control"),
             ("devgen", "config5_device_generator"), ("config4", "config4_year")):
    put("b_%s.json" % n, "bench_%s.json" % o)
for f in sorted(os.listdir(src)):   # multi-GPU lines, whatever the script named them
    if f.startswith("scale_") and f.endswith(".json"):
        put(f, f)

captures = []
for rep, out, phase, what in (("step_peak", "ncu_step_kernel_busy_hour.txt", "lockstep", "step 42 after reset: every pile occupied"),
                              ("step_quiet", "ncu_step_kernel_quiet_hour.txt", "lockstep", "step 22 after reset: a few cars per hub"),
                              ("step_spread", "ncu_step_kernel_dephased.txt", "spread", "de-phased batch: every time of day in one launch")):
    p = os.path.join(src, rep + ".ncu-rep")
    if not os.path.exists(p):
        continue
    txt = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_summary.py"), p, "45", "--regions"],
                         capture_output=True, text=True).stdout
    stalls = subprocess.run([sys.executable, os.path.join(root, "tools", "ncu_stalls.py"), p], capture_output=True, text=True).stdout
    open(os.path.join(dst, "%s_%s" % (pre, out)), "w").write(
        "# ncu --set full --clock-control none --import-source on, ONE chub_step_kernel launch over 524,288 default hubs (%s)\n%s\n%s"
        % (what, stalls, txt))
    raw = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, val = rows[0], rows[1], rows[2]

    def get(name):
        i = hdr.index(name)
        v = float(val[i].replace(",", ""))
        u = units[i].lower()
        return v * {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
    captures.append({"workload": "config5", "hubs": 524288, "phase": phase, "what": what,
                     "dram_bytes": int(get("dram__bytes_read.sum") + get("dram__bytes_write.sum")),
                     "duration_us_under_ncu": get("gpu__time_duration.sum"),
                     "warp_instructions": int(get("smsp__inst_executed.sum")),
                     "source": "profiles/%s_%s" % (pre, out)})
if captures:
    # bench.py quotes the capture whose phase matches its own: the busy-hour launch stands for lock step only in its
    # own file; the line of a lock-step bench takes the mean of the busy and the quiet launch as a plain label would lie
    lock = [c for c in captures if c["phase"] == "lockstep"]
    if len(lock) == 2:
        for c in lock:
            c["phase"] = "lockstep-" + ("busy" if "42" in c["what"] else "quiet")
    json.dump({"source": "ncu --set full --clock-control none, one launch each (tools/gpu_r02k.sh); dram__bytes_read.sum + "
                         "dram__bytes_write.sum", "captures": captures}, open(os.path.join(dst, "r02_traffic.json"), "w"), indent=1)
print("\n".join(sorted(f for f in os.listdir(dst) if f.startswith(pre))))
