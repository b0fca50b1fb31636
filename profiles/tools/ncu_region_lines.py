"""Per-region, per-source-line instruction breakdown of an `ncu --page source --print-source cuda,sass --csv` dump.
usage: ncu_region_lines.py dump.csv kernel_source.cuh REGION_NAME N marker1 marker2 ...
Regions are delimited (in SASS address order) by the first SASS instruction attributed to a kernel-source
line at or after each marker's line."""
import csv, collections, sys
fn, srcfile, want, topn = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
names = sys.argv[5:]
rows = list(csv.reader(open(fn)))
cur = hdr = curln = None
curtxt = ""
amap = {}
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; ix = {}
        for k, n in enumerate(hdr):
            ix.setdefault(n, k)
        continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        try: curln = int(r[0]); curtxt = r[1].strip()[:90]
        except ValueError: pass
        continue
    addr = r[2]
    if addr.startswith("0x"):
        try:
            amap[int(addr, 16)] = (cur, curln, curtxt, int(r[ix['# Samples']]), int(r[ix['Instructions Executed']]),
                                   int(r[ix['Thread Instructions Executed']]), r[3] if len(r) > 3 else "")
        except ValueError: pass
src = open(srcfile).read().splitlines()
marks = [(i, n) for i, l in enumerate(src, 1) for n in names if n in l]
base = srcfile.split('/')[-1]
def region(ln):
    r = "prologue"
    for i, n in marks:
        if ln >= i: r = n
    return r
agg = collections.defaultdict(lambda: [0, 0, 0, 0, ""])
curreg = "prologue"; tot = 0
for a in sorted(amap):
    f, ln, txt, smp, ins, thr, sass = amap[a]
    if f == base: curreg = region(ln)
    tot += ins
    if curreg == want:
        g = agg[(f, ln)]; g[0] += 1; g[1] += smp; g[2] += ins; g[3] += thr; g[4] = txt
rt = sum(g[2] for g in agg.values())
print("region %s: %d sass, %.1f%% of all warp-inst" % (want, sum(g[0] for g in agg.values()), 100 * rt / tot))
for (f, ln), g in sorted(agg.items(), key=lambda kv: -kv[1][2])[:topn]:
    print("%5.1f%% inst %4d sass thr/inst %5.1f | %s:%d | %s" % (100 * g[2] / rt, g[0], g[3] / max(g[2], 1), f, ln, g[4]))
