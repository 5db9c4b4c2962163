"""Per-opcode / per-region statistics from `ncu --page source --csv` output.
    ncu -i X.ncu-rep --page source --csv > src.csv ; python profiles/srcstat.py src.csv [regions]
"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr)]
num = lambda r, k: float(r[ix[k]] or 0)
ops = collections.Counter(); samp = collections.Counter(); wf = collections.Counter(); ideal = collections.Counter()
tot = 0
for r in data:
    src = r[ix["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDS", "STS", "LDG", "STG")) and "." in op else "")
    n = num(r, "Instructions Executed")
    ops[op] += n; tot += n
    samp[op] += num(r, "# Samples")
    wf[op] += num(r, "L1 Wavefronts Shared"); ideal[op] += num(r, "L1 Wavefronts Shared Ideal")
print("total warp-instructions executed: %.0f" % tot)
print("%-10s %12s %6s %9s %12s %12s" % ("opcode", "executed", "%", "samples%", "smem wf", "smem ideal"))
ts = sum(samp.values())
for op, n in ops.most_common(28):
    print("%-10s %12.0f %6.1f %9.1f %12.0f %12.0f" % (op, n, 100 * n / tot, 100 * samp[op] / max(ts, 1), wf[op], ideal[op]))
if len(sys.argv) > 2:
    # top stall instructions
    top = sorted(data, key=lambda r: -num(r, "# Samples"))[:40]
    for r in top:
        st = {k: num(r, k) for k in hdr if k.startswith("stall_") and "Not Issued" not in k}
        best = sorted(st.items(), key=lambda kv: -kv[1])[:2]
        print("%6.0f %s  %s  %s" % (num(r, "# Samples"), r[ix["Address"]][-5:], r[ix["Source"]][:70], best))
