"""Summarises gpurun_out/launches.csv (ncu launch list) and gpurun_out/prof.ncu-rep (ncu --set full) into
profiles/<tag>_launches.md and profiles/<tag>_full.md.   usage: python scripts/summarize_ncu.py <tag> [note]"""
import collections, csv, re, subprocess, sys

tag = sys.argv[1]
note = sys.argv[2] if len(sys.argv) > 2 else ""
rows = list(csv.DictReader(l for l in open("gpurun_out/launches.csv") if l.startswith('"')))
agg = collections.OrderedDict()
for r in rows:
    k = re.sub(r"\(.*", "", r["Kernel Name"].replace("unnamed>::", "").replace("void ", ""))
    a = agg.setdefault(k, [0, 0.0, r["Grid Size"], r["Block Size"]])
    a[0] += 1
    a[1] += float(r["Metric Value"])
tot = sum(v[1] for v in agg.values())
with open(f"profiles/{tag}_launches.md", "w") as f:
    f.write(f"# ncu launch list ({tag})\n\n{note}\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` over "
            "`python bench.py --steps 2 --warmup 1 --no-cpu-baseline` (cold-cache, serialised: compare shares).\n\n"
            "| kernel | launches | mean us | total ms | share | grid | block |\n|---|---|---|---|---|---|---|\n")
    for k, (n, t, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{k}` | {n} | {t/n/1e3:.1f} | {t/1e6:.3f} | {100*t/tot:.1f}% | {g} | {b} |\n")
raw = subprocess.run(["ncu", "-i", "gpurun_out/prof.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
if len(rr) > 2:
    hdr, units = rr[0], rr[1]
    idx = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum",
            "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
            "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"]
    with open(f"profiles/{tag}_full.md", "w") as f:
        f.write(f"# ncu --set full ({tag})\n\n{note}\n\n")
        for r in rr[2:]:
            f.write(f"## `{r[idx['Kernel Name']].replace('unnamed>::', '')[:90]}`\n\n| metric | value | unit |\n|---|---|---|\n")
            for w in want:
                if w in idx:
                    f.write(f"| {w} | {r[idx[w]]} | {units[idx[w]]} |\n")
            f.write("\n")
# dram bytes per launch of each kernel kind (default c2 workload) -> profiles/traffic.json, read by bench.py
import json, os
kind_of = [("ffn_fused_kernel", "ffn_fused"), ("qkv_attention_kernel", "qkv_attention"), ("attention_tc_kernel", "attention"), ("embed_layernorm", "embed_ln"),
           ("mean_pool", "pool"), ("gemm_tc_kernel<192, 0", "gemm_qkv"), ("gemm_tc_kernel<256, 1", "gemm_ffn1"),
           ("gemm_tc_kernel<192, 4", "gemm_attn_out")]
if len(rr) > 2:
    tj = "profiles/traffic.json"
    traffic = json.load(open(tj)) if os.path.exists(tj) else {}
    acc = {}
    for r in rr[2:]:
        name = r[idx["Kernel Name"]]
        for pat, kind in kind_of:
            if pat in name:
                scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
                b = float(r[idx["dram__bytes_read.sum"]]) * scale[units[idx["dram__bytes_read.sum"]]] + \
                    float(r[idx["dram__bytes_write.sum"]]) * scale[units[idx["dram__bytes_write.sum"]]]
                acc.setdefault(kind, []).append(b)
                break
    for kind, v in acc.items():
        traffic["c2:" + kind] = round(sum(v) / len(v))
    traffic["_source"] = f"profiles/{tag}_full.md (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch)"
    json.dump(traffic, open(tj, "w"), indent=1, sort_keys=True)
print("wrote profiles/%s_*.md" % tag)
