"""Markdown stall table of one kernel from `ncu -i rep --page source --csv` output (a CSV file or a .ncu-rep):
warp-state samples by reason, by code region (SASS lines grouped by execution count) and the top lines.
usage: python scripts/ncu_stall_table.py <source.csv|report.ncu-rep> <kernel substring> <out.md> [title]"""
import collections, csv, subprocess, sys
src, want, out = sys.argv[1], sys.argv[2], sys.argv[3]
title = sys.argv[4] if len(sys.argv) > 4 else ""
text = open(src).read() if src.endswith(".csv") else subprocess.run(["ncu", "-i", src, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks, cur = [], None
for r in csv.reader(text.splitlines()):
    if len(r) >= 2 and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
blk = next(b for b in blocks if want in b["name"])
hdr = blk["rows"][0]; idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in blk["rows"][1:] if len(r) == len(hdr)]
num = lambda r, k: int(r[idx[k]] or 0)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(num(r, "# Samples") for r in data)
with open(out, "w") as f:
    f.write(f"# warp-state samples: {title}\n\nkernel `{blk['name'][:110]}`\n\n`ncu --set full --import-source on`, source page; {tot} samples, {len(data)} SASS lines.\n\n")
    agg = {s: sum(num(r, s) for r in data) for s in stalls}
    f.write("| stall reason | samples | share |\n|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]:
        f.write(f"| {k[6:]} | {v} | {100*v/tot:.1f}% |\n")
    f.write("\n## by code region (SASS lines with the same execution count)\n\n| executed | lines | samples | top reasons |\n|---|---|---|---|\n")
    cnts = collections.Counter(num(r, "Instructions Executed") for r in data)
    for cnt, n in cnts.most_common(7):
        loop = [r for r in data if num(r, "Instructions Executed") == cnt]
        a2 = {s: sum(num(r, s) for r in loop) for s in stalls}
        f.write(f"| {cnt} | {n} | {sum(num(r, '# Samples') for r in loop)} | " + ", ".join(f"{k[6:]} {v}" for k, v in sorted(a2.items(), key=lambda kv: -kv[1])[:5] if v) + " |\n")
    f.write("\n## top SASS lines\n\n| samples | executed | instruction | reasons |\n|---|---|---|---|\n")
    for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:25]:
        st = sorted(((s[6:], num(r, s)) for s in stalls if num(r, s) > 0), key=lambda kv: -kv[1])[:3]
        f.write(f"| {num(r, '# Samples')} | {num(r, 'Instructions Executed')} | `{r[idx['Source']][:70]}` | {', '.join(f'{a} {b}' for a, b in st)} |\n")
print("wrote", out)
