"""Top source lines by one stall reason. usage: ncu_stall_lines.py dump.csv stall_long_sb [N]"""
import csv, sys, collections
fn, col = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
cur = hdr = None
agg = collections.defaultdict(lambda: [0, 0, ""])
for r in csv.reader(open(fn)):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; ix = {}
        for k, x in enumerate(hdr): ix.setdefault(x, k)
        continue
    if hdr is None or len(r) < len(hdr): continue
    try: ln = int(r[0]); v = int(r[ix[col]]); s = int(r[ix['# Samples']])
    except ValueError: continue
    a = agg[(cur, ln)]; a[0] += v; a[1] += s; a[2] = r[1].strip()[:100]
tot = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("%s: %d of %d samples (%.1f%%)" % (col, tot, ts, 100.0 * tot / ts))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    print("%5.1f%% | %s:%d | %s" % (100.0 * a[0] / tot, f, l, a[2]))
