"""Distance (in instructions) from each FP64 instruction of the sweep's hottest loop to the nearest
producer of one of its operands -- how well the rows are interleaved (development aid)."""
import collections, re, subprocess, sys
pat = sys.argv[1] if len(sys.argv) > 1 else "sweep_v2_kernelILi5ELb0ELb0"
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
for a, t in ins:
    m = re.search(r"BRA.*?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a:
        body = [(x, y) for x, y in ins if int(m.group(1), 16) <= x <= a]
        n = sum(1 for _, y in body if re.sub(r"^@!?U?P\w+\s+", "", y).startswith("DFMA"))
        if len(body) > 40 and n > 20 and (best is None or n / len(body) > best[0] / len(best[1])):
            best = (n, body)
body = best[1]
last, dist, nfp = {}, collections.Counter(), 0
for i, (_, t) in enumerate(body):
    t2 = re.sub(r"^@!?U?P\w+\s+", "", t)
    op = t2.split()[0]
    regs = [int(r) for r in re.findall(r"\bR(\d+)\b", t2)]
    if not regs:
        continue
    if op.startswith(("DFMA", "DMUL", "DADD")):
        nfp += 1
        dmin = min([i - last[r] for s_ in regs[1:] for r in (s_, s_ + 1) if r in last] or [99])
        dist[min(dmin, 8)] += 1
        last[regs[0]] = last[regs[0] + 1] = i
    else:
        last[regs[0]] = i
print(f"loop {len(body)} instr, {nfp} FP64;  producer distance histogram (8 = 8 or more):", dict(sorted(dist.items())))
