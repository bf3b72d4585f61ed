"""Here (no GPU): aggregate the `ncu --page source --csv` export of one kernel by SASS opcode: samples, executed
warp-instructions and the top stall reasons. usage: python tools/ncu_ops.py source.csv [N]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], rows[hi + 1:]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
seen, agg = set(), collections.defaultdict(lambda: [0, 0, collections.Counter()])


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


for d in data:
    if len(d) < len(hdr) or d[ix["Address"]] in seen or d[0] == "Address":
        continue
    seen.add(d[ix["Address"]])
    toks = [t for t in d[ix["Source"]].split() if not t.startswith("@")]
    op = toks[0] if toks else "?"
    if not op.startswith(("MUFU", "LDTM", "STTM", "SYNCS", "UTC", "UTMA")):
        op = op.split(".")[0]
    a = agg[op]
    a[0] += num(d[ix["# Samples"]])
    a[1] += num(d[ix["Instructions Executed"]])
    for s in stalls:
        v = num(d[ix[s]])
        if v:
            a[2][s] += v
tot = sum(a[0] for a in agg.values())
totex = sum(a[1] for a in agg.values())
print("samples", tot, "warp-instructions", totex)
for op, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print("%-34s samples %6d (%4.1f%%)  executed %10d (%4.1f%%)  %s" % (op, a[0], 100 * a[0] / tot, a[1], 100 * a[1] / totex,
                                                                       [(k[6:], v) for k, v in a[2].most_common(3)]))
