#!/usr/bin/env python
"""Dynamic SASS opcode histogram + stall reasons from `ncu -i X.ncu-rep --page source --csv
--print-source sass` output (one or more kernels per file)."""
import collections
import csv
import re
import sys


def main(path, top=28):
    rows = list(csv.reader(open(path)))
    kernels = []
    cur = None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            kernels.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None:
            cur["rows"].append(r)
    for k in kernels:
        hdr = k["hdr"]
        ia, ie, it, iss = (hdr.index(x) for x in ("Source", "Instructions Executed",
                                                   "Thread Instructions Executed", "# Samples"))
        stallcols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
        tot, samp, stalls = collections.Counter(), collections.Counter(), collections.Counter()
        wi = thr = 0
        for r in k["rows"]:
            if len(r) <= ie or not r[ie].isdigit():
                continue
            src = re.sub(r"^@!?U?P\w+\s+", "", r[ia].strip())
            op = src.split()[0] if src else "?"
            m = re.match(r"(IMAD\.(MOV|IADD|SHL|WIDE|HI)|[A-Z0-9]+)", op)
            op = m.group(1) if m else op
            n = int(r[ie])
            tot[op] += n
            wi += n
            thr += int(r[it])
            samp[op] += int(r[iss] or 0)
            for i in stallcols:
                if r[i].isdigit():
                    stalls[hdr[i]] += int(r[i])
        print(f"== {k['name']}: warp instr {wi}, thread instr {thr}, avg active {thr / max(wi, 1):.1f}")
        for op, n in tot.most_common(top):
            print(f"  {op:12s} {n:10d} {100 * n / wi:5.1f}%  samples {samp[op]}")
        print("  stalls:", stalls.most_common(10))


if __name__ == "__main__":
    main(sys.argv[1])
