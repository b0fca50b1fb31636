"""Static code size per source line / file for one kernel of an `nvdisasm -g -c` dump.
usage: sass_size_by_line.py all.dis KERNEL_SUBSTRING [N] [ADDR_LO ADDR_HI]"""
import re, sys, collections
fn, key = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
lo = int(sys.argv[4], 16) if len(sys.argv) > 5 else 0
hi = int(sys.argv[5], 16) if len(sys.argv) > 5 else 1 << 40
inside = False
cur = ("?", 0)
by_line = collections.Counter(); by_file = collections.Counter(); total = 0
for l in open(fn):
    if l.startswith(".text."):
        inside = key in l
        continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,6})\*/\s+\S', l)
    if m:
        a = int(m.group(1), 16)
        if lo <= a < hi:
            by_line[cur] += 1; by_file[cur[0]] += 1; total += 1
print("instructions: %d (%.1f KB)" % (total, total * 16 / 1024.0))
for f, c in by_file.most_common(): print("  %-28s %5d  %5.1f KB" % (f, c, c * 16 / 1024.0))
for (f, ln), c in by_line.most_common(n): print("  %4d  %s:%d" % (c, f, ln))
