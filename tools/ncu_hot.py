"""Top stall sites of one kernel from `ncu --page source --csv` output. usage: ncu_hot.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr_idx = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for k, hi in enumerate(hdr_idx[:1]):
    h = rows[hi]
    end = hdr_idx[k + 1] - 1 if k + 1 < len(hdr_idx) else len(rows)
    isamp, isrc = h.index("# Samples"), h.index("Source")
    stall = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    data = [r for r in rows[hi + 1:end] if len(r) == len(h) and r[isamp].isdigit()]
    tot = sum(int(r[isamp]) for r in data) or 1
    agg = {}
    for r in data:
        for i in stall:
            agg[h[i]] = agg.get(h[i], 0) + int(r[i] or 0)
    print("samples", tot, sorted(((v, n) for n, v in agg.items()), reverse=True)[:6])
    data.sort(key=lambda r: -int(r[isamp]))
    for r in data[:top]:
        st = sorted(((int(r[i] or 0), h[i][6:]) for i in stall), reverse=True)[:2]
        print("%5.1f%% %-16s %-84s %s" % (100 * int(r[isamp]) / tot, r[0][-8:], r[isrc][:84], st))
