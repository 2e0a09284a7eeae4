#!/usr/bin/env python
"""Static opcode mix of one kernel in a cuobjdump -sass listing (the S-box epilogues are straight-line code, so the static mix of the
large basic blocks tracks the executed mix ncu reports).  usage: sass_mix.py file.sass <substring of the mangled name> [substring ...]"""
import collections, re, sys
txt = open(sys.argv[1]).read().split("Function : ")
for want in sys.argv[2:]:
    for blk in txt[1:]:
        name = blk.split("\n", 1)[0]
        if want not in name:
            continue
        ops = collections.Counter()
        for line in blk.splitlines():
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if m:
                ops[".".join(m.group(1).split(".")[:2])] += 1
        tot = sum(ops.values())
        fp = ops["DFMA.RZ"] + ops["DADD"] + ops["DFMA"] + ops["DMUL"]
        alu = sum(v for k, v in ops.items() if k.split(".")[0] in ("IADD3", "LOP3", "SHF", "LEA", "ISETP", "SEL", "VIADD", "PRMT", "IABS", "IMNMX", "BMSK", "SGXT", "FLO", "POPC", "PLOP3"))
        print("%s: %d instructions, FP64 %d, ALU-pipe %d" % (name, tot, fp, alu))
        print("   " + "  ".join("%s %d" % kv for kv in ops.most_common(16)))
