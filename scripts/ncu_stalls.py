"""Summarise the source page of an ncu report: stall-reason totals and the top SASS lines by samples.
usage: python scripts/ncu_stalls.py <report.ncu-rep> [top_n]"""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
num = lambda r, k: int(r[idx[k]] or 0)
tot = sum(num(r, "# Samples") for r in data)
print("kernel:", rows[0][1][:120]); print("total samples", tot, "SASS lines", len(data))
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print(", ".join(f"{k[6:]} {v} ({100*v/tot:.0f}%)" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]))
for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:topn]:
    st = sorted(((s[6:], num(r, s)) for s in stalls if num(r, s) > 0), key=lambda kv: -kv[1])[:3]
    print(r[idx["Address"]][-5:], f'{num(r, "# Samples"):6d}', f'{num(r, "Instructions Executed"):9d}', r[idx["Source"]][:64].ljust(64), st)
