"""Per-source-line executed warp instructions of the first launch in an .ncu-rep, for a file and line range.
usage: ncu_lines.py rep file_prefix first last [warps]   (counts are divided by `warps` when given)"""
import csv, subprocess, sys, io
rep, pref, a, b = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
warps = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, seen, skip = None, set(), False
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        skip = cur in seen
        seen.add(cur)
        continue
    if skip or cur is None or not cur.startswith(pref) or r[0] in ("Function Name", "Line No", ""):
        continue
    try:
        ln, n, t = int(r[0]), int(r[7]), int(r[8])
    except Exception:
        continue
    if a <= ln <= b and n:
        print("%5d %9.1f lanes %4.1f | %s" % (ln, n / warps, t / n, r[1].strip()[:120]))
