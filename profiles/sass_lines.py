"""Join an ncu `--page source --csv` SASS export with nvdisasm -g line info:
samples per source line for one kernel.
usage: sass_lines.py <nvdisasm -g -c output> <ncu source csv> <kernel substring> [top]"""
import collections
import csv
import re
import sys

sass, ncucsv, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# address -> (file, line) from nvdisasm
addr2line = {}
cur = None
infunc = False
for ln in open(sass, errors="replace"):
    if ln.startswith(".text.") or re.match(r"\s*\.section\s+\.text\.", ln):
        infunc = kern in ln
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and infunc:
        addr2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(ncucsv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
iS = hdr.index("# Samples")
iI = hdr.index("Instructions Executed")
base = None
per = collections.Counter()
inst = collections.Counter()
for r in rows[hi + 1:]:
    try:
        a = int(r[0], 16) if r[0].startswith("0x") else int(r[0])
    except ValueError:
        continue
    if base is None:
        base = a
    key = addr2line.get(a - base)
    per[key] += int(r[iS] or 0)
    inst[key] += int(r[iI] or 0)
tot = sum(per.values())
print("total samples", tot, "total warp-instructions", sum(inst.values()))
for k, v in per.most_common(top):
    print("%6.2f%%  %9d inst  %s" % (100.0 * v / tot, inst[k], k))
