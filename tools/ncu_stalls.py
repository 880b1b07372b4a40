"""Headline metrics + warp-stall breakdown of each launch in an .ncu-rep (ncu --set full)."""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
def col(name):
    return hdr.index(name) if name in hdr else -1
want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__cycles_active.avg', 'sm__cycles_active.max', 'sm__cycles_elapsed.max', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__waves_per_multiprocessor', 'smsp__warps_eligible.avg.per_cycle_active',
        'sm__inst_executed.avg.per_cycle_elapsed']
for r in rows[2:]:
    print("==", r[col('Kernel Name')][:60] if col('Kernel Name') >= 0 else "")
    for w in want:
        if col(w) >= 0:
            print("  %-66s %s %s" % (w, r[col(w)], rows[1][col(w)]))
    st = [(float(r[i].replace(",", "")), h) for i, h in enumerate(hdr)
          if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and r[i]]
    if not st:
        st = [(float(r[i].replace(",", "")), h) for i, h in enumerate(hdr)
              if h.startswith("smsp__average_warp_latency_issue_stalled_") and r[i]]
    tot = sum(x[0] for x in st) or 1
    for v, h in sorted(st, reverse=True)[:12]:
        print("  stall %-58s %6.2f (%4.1f%%)" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v, 100 * v / tot))
