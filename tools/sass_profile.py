"""Opcode histogram weighted by executed warp instructions, and the hottest SASS lines, from `ncu --page source --csv
--print-source sass` (first kernel in the file)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iE, iT, iSt = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
ops, ops_t, tot, samples = collections.Counter(), collections.Counter(), 0, collections.Counter()
lines = []
for r in rows[2:]:
    if len(r) != len(hdr):
        break   # next kernel
    src = r[iS].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    op = op.split(".")[0]
    e, t, s = int(r[iE] or 0), int(r[iT] or 0), int(r[iSt] or 0)
    ops[op] += e; ops_t[op] += t; tot += e; samples[op] += s
    lines.append((e, s, src))
print("total warp instructions", tot)
for op, e in ops.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print("%-10s %12d  %5.1f%%  avg threads %4.1f  stall samples %d" % (op, e, 100.0 * e / tot, ops_t[op] / max(e, 1), samples[op]))
