import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) >= len(hdr)]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
seg, segs = [], []
for r in data:
    seg.append(r)
    if "BAR.SYNC" in r[ix["Source"]] or "EXIT" in r[ix["Source"]]:
        segs.append(seg); seg = []
if seg: segs.append(seg)
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
toti = sum(int(r[ix["Instructions Executed"]] or 0) for r in data)
print("total samples", tot, "total inst %.2fM" % (toti / 1e6), "segments", len(segs))
for k, sg in enumerate(segs):
    smp = sum(int(r[ix["# Samples"]] or 0) for r in sg)
    inst = sum(int(r[ix["Instructions Executed"]] or 0) for r in sg)
    if smp < tot * 0.01: continue
    st = {c: sum(int(r[ix[c]] or 0) for r in sg) for c in stall_cols}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:6]
    ops = {}
    for r in sg:
        t = r[ix["Source"]].split()
        op = (t[1] if t and t[0].startswith("@") else (t[0] if t else "?"))
        ops[op] = ops.get(op, 0) + int(r[ix["Instructions Executed"]] or 0)
    topops = sorted(ops.items(), key=lambda kv: -kv[1])[:8]
    print(f"seg {k:2d} sass {len(sg):4d} inst {inst/1e6:6.2f}M ({100*inst/toti:4.1f}%) samples {smp:6d} ({100*smp/tot:4.1f}%)", [(a.replace('stall_',''), b) for a, b in top])
    print("        ", [(a, round(b/1e6, 2)) for a, b in topops])
