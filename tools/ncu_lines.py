"""Aggregate an `ncu --page source --print-source cuda,sass --csv` export by source line."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
cur = None; agg = collections.defaultdict(lambda: [0.0, 0.0, ""])
hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; ie = hdr.index("Instructions Executed"); sm = hdr.index("# Samples"); continue
    if hdr is None or len(r) <= ie: continue
    if r[2] != "-": continue           # SASS rows carry an address; source rows have "-"
    try: a = float(r[ie]); s = float(r[sm])
    except ValueError: continue
    k = (cur, int(r[0])); agg[k][0] += a; agg[k][1] += s; agg[k][2] = r[1].strip()
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print("total warp-inst %d samples %d" % (ti, ts))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print("%5.1f%% inst %5.1f%% smp  %s:%d  %s" % (100 * v[0] / ti, 100 * v[1] / ts, k[0], k[1], v[2][:100]))
