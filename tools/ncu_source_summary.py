"""Summarise `ncu --page source --csv` output: stall reasons overall, samples per opcode class and
the hottest SASS regions.   python tools/ncu_source_summary.py file.csv [top]"""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]
col = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = collections.Counter(); byop = collections.Counter(); nsamp = 0; execs = collections.Counter()
data = []
for r in rows[hi + 1:]:
    if len(r) < len(h): continue
    s = int(r[col["# Samples"]] or 0)
    src = r[col["Source"]].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    op = op.split(".")[0]
    nsamp += s; byop[op] += s; execs[op] += int(r[col["Instructions Executed"]] or 0)
    for n in stalls: tot[n] += int(r[col[n]] or 0)
    data.append((s, src, {n: int(r[col[n]] or 0) for n in stalls}))
print("total samples", nsamp)
print("stall reasons:", ", ".join(f"{k[6:]}={v / max(nsamp,1):.3f}" for k, v in tot.most_common(10)))
print("samples by opcode:", ", ".join(f"{k}={v / max(nsamp,1):.3f}" for k, v in byop.most_common(14)))
print("executed by opcode:", ", ".join(f"{k}={v}" for k, v in execs.most_common(14)))
print("hottest instructions:")
idx = sorted(range(len(data)), key=lambda i: -data[i][0])[:top]
for i in sorted(idx):
    s, src, st = data[i]
    main = max(st.items(), key=lambda kv: kv[1])
    print(f"  #{i:5d} {s:7d} {s / nsamp:.3f}  {src[:70]:70s} {main[0][6:]}")
# cumulative profile in 20 equal chunks of the instruction stream
n = len(data); ch = max(n // 20, 1)
print("samples per 5% chunk of the SASS stream:", [round(sum(d[0] for d in data[k:k + ch]) / nsamp, 3) for k in range(0, n, ch)])
