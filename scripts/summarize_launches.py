"""Summarise an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` launch list of
`bench.py --ncu-mode`: per-kernel time share and DRAM bytes of the last complete decode step (sampler to sampler),
and of the prompt pass. Writes a markdown table and profiles/traffic.json.
usage: python scripts/summarize_launches.py gpurun_out/r01_bench_launches.csv profiles/r01_decode_step.md"""
import collections, csv, json, os, sys

src, out = sys.argv[1], sys.argv[2]
with open(src) as f:
    rows = list(csv.DictReader([l for l in f if not l.startswith("==")]))
L = collections.OrderedDict()
for r in rows:
    d = L.setdefault(int(r["ID"]), {"name": r["Kernel Name"], "grid": r["Grid Size"]})
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]; m = r["Metric Name"]
    if m == "gpu__time_duration.sum":
        d["us"] = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    else:
        d[m] = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
ids = list(L)


def kind(d):
    n = d["name"]
    if "matvec_kernel" in n:
        g = int(d["grid"].strip("()").split(",")[0])
        return "matvec_kernel"
    return n.split("(")[0].replace("void ", "").replace("b2::", "")


def table(sel, title):
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for d in sel:
        k = kind(d)
        agg[k][0] += 1; agg[k][1] += d["us"]; agg[k][2] += d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
    tot = sum(v[1] for v in agg.values()); tb = sum(v[2] for v in agg.values())
    lines = [f"### {title}", "", "| kernel | launches | time (us, serialised under ncu) | share | DRAM read+write (MB) |", "|---|---|---|---|---|"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| `{k}` | {v[0]} | {v[1]:.1f} | {100 * v[1] / tot:.1f} % | {v[2] / 1e6:.1f} |")
    lines.append(f"| **total** | {len(sel)} | {tot:.1f} | 100 % | {tb / 1e6:.1f} |")
    return "\n".join(lines), tot, tb


samp = [i for i in ids if "sample_kernel" in L[i]["name"]]
a, b = samp[-2] + 1, samp[-1] + 1
step = [L[i] for i in ids if a <= i < b]
t1, tot, tb = table(step, "One decode step (last complete step of the run: 163 launches = one CUDA-graph replay)")
# prompt pass: from the second embed_rows launch (warm prefill) to the prefill_finish that follows it
emb = [i for i in ids if "embed_rows_kernel" in L[i]["name"]]
fin = [i for i in ids if "prefill_finish_kernel" in L[i]["name"]]
parts = [t1]
if len(emb) >= 2 and len(fin) >= 2:
    pre = [L[i] for i in ids if emb[1] <= i <= fin[1]]
    t2, _, _ = table(pre, "One 512-token prompt pass (second, warm b2_prefill)")
    parts.append(t2)
open(out, "w").write("\n\n".join(parts) + "\n")
json.dump({"dram_bytes_per_step": int(tb), "launches_per_step": len(step), "serialised_us_per_step": round(tot, 1),
           "source": os.path.basename(src) + " (ncu dram__bytes_read.sum + dram__bytes_write.sum over the 163 launches of one decode step)"},
          open(os.path.join(os.path.dirname(out), "traffic.json"), "w"), indent=1)
print(open(out).read())
