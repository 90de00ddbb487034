"""Summarise the source page of an ncu report: stall-reason totals and the top SASS lines by samples.
usage: python scripts/ncu_stalls.py <report.ncu-rep> [top_n] [kernel-name substring]"""
import csv, subprocess, sys, collections
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
want = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# the page repeats ["Kernel Name", name] + header + lines for every captured launch
blocks, cur = [], None
for r in rows:
    if len(r) >= 2 and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
blk = next(b for b in blocks if want in b["name"])
hdr = blk["rows"][0]; idx = {h: i for i, h in enumerate(hdr)}; data = [r for r in blk["rows"][1:] if len(r) == len(hdr)]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
num = lambda r, k: int(r[idx[k]] or 0)
tot = sum(num(r, "# Samples") for r in data)
print("kernel:", blk["name"][:120]); print("total samples", tot, "SASS lines", len(data))
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print(", ".join(f"{k[6:]} {v} ({100*v/tot:.0f}%)" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]))
cnts = collections.Counter(num(r, "Instructions Executed") for r in data)
for cnt, n in cnts.most_common(6):
    loop = [r for r in data if num(r, "Instructions Executed") == cnt]
    a2 = {s: sum(num(r, s) for r in loop) for s in stalls}
    print(f"  executed {cnt:9d} x: {n:4d} SASS lines, {sum(num(r, '# Samples') for r in loop):6d} samples:",
          ", ".join(f"{k[6:]} {v}" for k, v in sorted(a2.items(), key=lambda kv: -kv[1])[:6] if v))
for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:topn]:
    st = sorted(((s[6:], num(r, s)) for s in stalls if num(r, s) > 0), key=lambda kv: -kv[1])[:3]
    print(r[idx["Address"]][-5:], f'{num(r, "# Samples"):6d}', f'{num(r, "Instructions Executed"):9d}', r[idx["Source"]][:64].ljust(64), st)
