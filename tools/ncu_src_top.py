"""Developer tool: top stall lines of an `ncu --page source --csv` export (one kernel)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = [k for k, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
col = {h: i for i, h in enumerate(hdr)}
iS, isrc, iex, iw, iexc = col["Warp Stall Sampling (All Samples)"], col["Source"], col["Instructions Executed"], col["L1 Wavefronts Shared"], col["L1 Wavefronts Shared Excessive"]
tot = sum(int(r[iS] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
print("excess shared wavefronts by opcode:")
agg = {}
for r in data:
    op = r[isrc].split()[0] if r[isrc].split() else ""
    if op.startswith("@"):
        op = r[isrc].split()[1]
    e = int(r[iexc] or 0)
    w = int(r[iw] or 0)
    a = agg.setdefault(op, [0, 0, 0])
    a[0] += e; a[1] += w; a[2] += int(r[iex] or 0)
for op, (e, w, x) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
    print("  %-28s wavefronts %12d excessive %12d executed %12d" % (op, w, e, x))
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
top = sorted(range(len(data)), key=lambda k: -int(data[k][iS] or 0))[:n]
for k in sorted(top):
    r = data[k]
    st = sorted(((int(r[col[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:2]
    print("%5d %6s ex=%-9s wf=%-9s exc=%-8s %-70s %s" % (k, r[iS], r[iex], r[iw], r[iexc], r[isrc][:70], st))
