"""Static code size of one kernel by source region: nvdisasm -g output -> SASS instructions per "@region" (the same markers
tools/ncu_summary.py uses for the dynamic counts).  Usage: python tools/sass_regions.py all.sass '<mangled kernel name>'"""
import collections, glob, os, re, sys
sass, kern = sys.argv[1], sys.argv[2]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REG = []
for path in glob.glob(os.path.join(root, "charginghub_env_b200", "csrc", "*")) + glob.glob(os.path.join(root, "include", "*.h")):
    if not path.endswith((".cu", ".cuh", ".h")):
        continue
    marks = [(k + 1, m.group(1).strip()) for k, line in enumerate(open(path, errors="ignore"))
             for m in [re.search(r"@region\s+(.*)", line)] if m]
    base = os.path.basename(path)
    if not marks:
        REG.append((base, 0, 10 ** 9, base))
    for j, (ln, lab) in enumerate(marks):
        REG.append((base, ln, marks[j + 1][0] - 1 if j + 1 < len(marks) else 10 ** 9, lab))
inside, cur = False, ("?", 0)
cnt, lines = collections.Counter(), collections.Counter()
for line in open(sass, errors="ignore"):
    if line.startswith(".text."):
        inside = line.strip().rstrip(":") == ".text." + kern
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        lab = "other " + cur[0]
        for f, a, b, l in REG:
            if f == cur[0] and a <= cur[1] <= b:
                lab = l
                break
        cnt[lab] += 1
        lines[cur] += 1
tot = sum(cnt.values())
print("SASS instructions:", tot, "(%.0f KB)" % (tot * 16 / 1024))
for lab, n in cnt.most_common():
    print("%6d %5.1f%%  %s" % (n, 100.0 * n / tot, lab))
print("-- heaviest source lines")
for (f, ln), n in lines.most_common(25):
    print("%6d  %s:%d" % (n, f, ln))
