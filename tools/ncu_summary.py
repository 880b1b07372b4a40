"""Summarise an .ncu-rep: headline metrics per launch and the hottest source lines (needs -lineinfo)."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'sm__inst_executed_pipe_fp64.sum',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_lsu.sum',
        'sm__inst_executed_pipe_xu.sum', 'smsp__inst_executed_op_branch.sum',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_srcunit_tex_op_write.sum']
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print("%-70s %-12s %s" % (w, rows[1][i], [r[i][:40] for r in rows[2:]]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
recs, seen, cur, skip = [], set(), None, False
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        skip = cur in seen
        seen.add(cur)
        continue
    if skip or r[0] in ("Function Name", "Line No") or r[0] == "":
        continue
    try:
        recs.append((int(r[7]), int(r[4]), int(r[8]), cur, int(r[0]), r[1].strip()[:100]))
    except Exception:
        pass
tot = sum(x[0] for x in recs) or 1
tots = sum(x[1] for x in recs) or 1
print("warp-inst (first launch)", tot, "stall samples", tots)
for n, w, t, f, ln, text in sorted(recs, key=lambda x: -x[0])[:topn]:
    print("%5.2f%% inst %5.2f%% stall lanes %4.1f | %s:%d %s" % (100.0 * n / tot, 100.0 * w / tots, (t / n) if n else 0, f[:14], ln, text))

# ---- per-region totals.  Regions are declared in the sources themselves: a comment "@region <label>" starts a
# region that runs to the next marker of that file.  Lines of inlined device functions are attributed to their
# own file:line, whatever the caller.
import glob, os, re
REGIONS = []
_csrc = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
for path in glob.glob(os.path.join(_csrc, "charginghub_env_b200", "csrc", "*")) + glob.glob(os.path.join(_csrc, "include", "*.h")):
    if not path.endswith((".cu", ".cuh", ".h")):
        continue
    marks = [(k + 1, m.group(1).strip()) for k, line in enumerate(open(path, errors="ignore"))
             for m in [re.search(r"@region\s+(.*)", line)] if m]
    base = os.path.basename(path)
    if not marks:
        REGIONS.append((base, 0, 10 ** 9, base))
    for j, (ln, lab) in enumerate(marks):
        REGIONS.append((base, ln, marks[j + 1][0] - 1 if j + 1 < len(marks) else 10 ** 9, lab))
if "--regions" in sys.argv:
    agg = collections.OrderedDict()
    for n, w, t, f, ln, text in recs:
        lab = "other " + f
        for rf, a, b, l in REGIONS:
            if rf.startswith(f[:14]) and f[:14] == rf[:14] and a <= ln <= b:
                lab = l
                break
        x = agg.setdefault(lab, [0, 0, 0])
        x[0] += n; x[1] += w; x[2] += t
    print("region totals (warp-inst share, stall share, active lanes)")
    for lab, (n, w, t) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print("%6.2f%% inst %6.2f%% stall lanes %4.1f | %s" % (100.0 * n / tot, 100.0 * w / tots, (t / n) if n else 0, lab))
