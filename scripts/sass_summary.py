"""Counts the Blackwell-specific SASS mnemonics per kernel of libb200_embed.so (cuobjdump -sass): the evidence that the hot
kernels issue tcgen05 MMAs (UTCHMMA), tensor-memory loads / stores (LDTM / STTM) and TMA copies (UTMALDG / UTMASTG).
usage: python scripts/sass_summary.py > profiles/r02_sass_summary.md"""
import collections, os, re, subprocess
lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "codename-symbiont_b200", "libb200_embed.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMALDG.2D.MULTICAST", "UTMASTG", "UTCBAR", "SYNCS", "MUFU.EX2", "HMMA"]
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "").replace("void ", ""))
        cur = counts.setdefault(name, collections.Counter())
        continue
    if cur is None:
        continue
    for p in pats:
        if re.search(r"\b" + re.escape(p) + r"(\.|\b)", line):
            cur[p] += 1
print("# SASS mnemonics per kernel (`cuobjdump -sass codename-symbiont_b200/libb200_embed.so`, sm_100a)\n")
print("Counts of instructions per kernel instantiation; kernels without any of them (elementwise, search, fp32 validation) are listed at the end.\n")
print("| kernel | " + " | ".join(pats) + " |\n|---|" + "---|" * len(pats))
plain = []
for name, c in counts.items():
    if not any(c[p] for p in pats if p != "MUFU.EX2" and p != "SYNCS"):
        plain.append(name)
        continue
    print(f"| `{name[:110]}` | " + " | ".join(str(c[p]) for p in pats) + " |")
print("\nWithout tensor-core / TMA instructions: " + ", ".join(f"`{n[:60]}`" for n in sorted(set(plain))))
