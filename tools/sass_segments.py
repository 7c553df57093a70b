"""Groups the SASS lines of an `ncu --page source --csv --print-source sass` export into runs of equal execution count:
shows how many instructions each loop body has, how often it ran and with how many active threads."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iE, iT, iSt = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
out = []
for k, r in enumerate(rows[2:]):
    if len(r) != len(hdr):
        break
    out.append((k, int(r[iE] or 0), int(r[iT] or 0), int(r[iSt] or 0), r[iS].strip()))
tot = sum(o[1] for o in out)
minshare = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
start, acc, tacc, samp, first, prev = 0, 0, 0, 0, "", None
def flush(end):
    if prev is not None and 100.0 * acc / tot >= minshare:
        print("%4d-%4d n=%3d exec=%9d thr=%4.1f share=%4.1f%% samples=%6d | %s" % (start, end, end - start + 1, prev, tacc / max(acc, 1), 100.0 * acc / tot, samp, first))
for k, e, t, s, src in out:
    if prev is None or abs(e - prev) > 0.02 * max(prev, 1):
        flush(k - 1)
        start, acc, tacc, samp, first = k, 0, 0, 0, src[:50]
    prev = e; acc += e; tacc += t; samp += s
flush(out[-1][0])
print("total warp instructions", tot)
