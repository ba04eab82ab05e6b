"""Samples of one kernel in an .ncu-rep, summed per source line (development aid).
usage: python tools/ncu_lines.py <report.ncu-rep> <mangled-name substring> [n lines]
Needs the library built from the same sources (line info comes from nvdisasm on the in-tree .so)."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(root, "melissa_b200", "libmelissa_b200.so")], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(txt) if l.startswith(".text.") and pat in l)
line, addr2line = None, {}
for l in txt[start + 1:]:
    if (l.startswith(".text.") or l.startswith(".section")) and addr2line:
        break
    m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', l)
    if m:
        line = (m.group(1), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m:
        addr2line[int(m.group(1), 16)] = line
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ia, ismp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
cols = {h: i for i, h in enumerate(hdr)}
base = int(data[0][ia], 16)
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot = 0
kinds = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in data:
    ln = addr2line.get(int(r[ia], 16) - base)
    s = int(r[ismp] or 0)
    tot += s
    e = agg[ln]
    e[0] += s
    e[1] += int(r[iex] or 0)
    for k in kinds:
        e[2][k] += int(r[cols[k]] or 0)
src = {}
for f in os.listdir(os.path.join(root, "melissa_b200", "csrc")):
    src[f] = open(os.path.join(root, "melissa_b200", "csrc", f)).read().splitlines()
print("total samples", tot)
allk = collections.Counter()
for _, (_, _, c) in agg.items():
    allk.update(c)
print("by reason:", ", ".join("%s=%.1f%%" % (k[6:], 100 * v / tot) for k, v in allk.most_common(8)))
for ln, (s, e, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    text = src.get(ln[0], [""])[ln[1] - 1].strip()[:90] if ln and ln[0] in src and ln[1] - 1 < len(src[ln[0]]) else ""
    print("%5.1f%% %-11s:%-4s ex=%-11d %-28s %s" % (100 * s / tot, ln[0] if ln else None, ln[1] if ln else "", e,
                                                   ",".join("%s=%d" % (k[6:], v) for k, v in c.most_common(2)), text))
