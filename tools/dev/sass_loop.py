"""Opcode mix of the hottest loop (the backward-branch range with most DFMA) of one kernel (development aid).
usage: python tools/dev/sass_loop.py <substring of the mangled kernel name> [rows per trip]"""
import collections
import re
import subprocess
import sys

pat = sys.argv[1]
rows = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
txt = subprocess.run(["cuobjdump", "-sass", "melissa_b200/libmelissa_b200.so"], capture_output=True, text=True).stdout
ins, on = [], False
for line in txt.splitlines():
    if "Function :" in line:
        on = pat in line
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if on and m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
best = None
for addr, text in ins:
    m = re.search(r"BRA(?:\.\w+)*\s+(?:\w+,\s*)?`?\(?\.?L?_?x?_?(\w+)\)?", text)
    m2 = re.search(r"BRA.*?0x([0-9a-f]+)", text)
    if m2:
        tgt = int(m2.group(1), 16)
        if tgt < addr:
            body = [t for a, t in ins if tgt <= a <= addr]
            n = sum(1 for t in body if re.sub(r"^@!?U?P\w+\s+", "", t).startswith("DFMA"))
            if len(body) > 40 and n > 20 and (best is None or n / len(body) > best[0] / len(best[3])):
                best = (n, tgt, addr, body)
if best is None:
    sys.exit("no loop found")
n, tgt, addr, body = best
c = collections.Counter(re.sub(r"^@!?U?P\w+\s+", "", t).split()[0].split(".")[0] for t in body)
print(f"loop 0x{tgt:x}..0x{addr:x}: {len(body)} instr, per row {len(body) / rows:.1f}")
fp64 = sum(v for k, v in c.items() if k in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"))
print(f"  FP64 {fp64 / rows:.1f}  other {(len(body) - fp64) / rows:.1f}")
print("  " + "  ".join(f"{k}:{v / rows:.1f}" for k, v in c.most_common(24)))
