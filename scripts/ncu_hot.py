"""Hot SASS regions of one kernel from `ncu --page source --csv` output: instruction share + stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.006
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
data = [r for r in rows[h + 1:] if len(r) > 10]
# the page lists the SASS twice when source correlation is on: keep the first copy
first = data[0][0]
dup = [i for i, r in enumerate(data) if r[0] == first]
if len(dup) > 1:
    data = data[:dup[1]]
ie, ss, src = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
num = lambda s: int(s) if s.isdigit() else 0
tot, tots = sum(num(r[ie]) for r in data), sum(num(r[ss]) for r in data)
print("warp instructions", tot, "samples", tots, "sass rows", len(data))
for k, r in enumerate(data):
    n, s = num(r[ie]), num(r[ss])
    if n >= thr * tot or s >= 2 * thr * tots:
        print("%5d inst %5.2f%%  smp %5.2f%%  %s" % (k, 100 * n / tot, 100 * s / tots, r[src].strip()[:95]))
if len(sys.argv) > 3:
    w = int(sys.argv[3])
    print("--- regions of %d SASS rows" % w)
    for k in range(0, len(data), w):
        n = sum(num(r[ie]) for r in data[k:k + w]); s = sum(num(r[ss]) for r in data[k:k + w])
        if n > 0.01 * tot or s > 0.01 * tots:
            print("%5d-%5d inst %5.2f%%  smp %5.2f%%   %s" % (k, k + w - 1, 100 * n / tot, 100 * s / tots, data[k][src].strip()[:60]))
