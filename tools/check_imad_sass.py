"""Prints the instruction mix of the timed loop of every mode of a micro-benchmark binary (largest backward-branch body)."""
import re, collections, subprocess, sys
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
name = None; lines = collections.defaultdict(list)
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m: name = m.group(1); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and name: lines[name].append((int(m.group(1), 16), m.group(2).strip()))
for n, ins in sorted(lines.items()):
    best = None
    for addr, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?0x([0-9a-f]+)", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < addr and (best is None or addr - tgt > best[1] - best[0]): best = (tgt, addr)
    body = [t for a, t in ins if best and best[0] <= a <= best[1]]
    c = collections.Counter()
    for t in body:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        op = t.split()[0]
        if op == "IMAD.WIDE.U32": op += "(RZ)" if t.rstrip().endswith("RZ") else "(acc)"
        c[op] += 1
    print(n, len(body), dict(c.most_common(10)))
