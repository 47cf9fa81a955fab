"""Hot SASS instructions of an .ncu-rep source page: python tools/ncu_hot.py rep [topN]  (reads `ncu --page source --csv`)"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; body = []
for r in rows[hi + 1:]:   # first kernel of the report only
    if r and r[0] in ("Address", "Kernel Name"): break
    if len(r) == len(hdr): body.append(r)
c = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[c["# Samples"]] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
order = sorted(range(len(body)), key=lambda i: -int(body[i][c["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[c[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:3]
    print(f"{i:5d} {int(r[c['# Samples']]):7d} {100*int(r[c['# Samples']])/tot:5.1f}%  exec={r[c['Instructions Executed']]:>10}  {r[c['Source']].strip()[:90]:90s} {[f'{n}:{v}' for v, n in st if v]}")
