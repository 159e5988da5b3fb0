"""Turn gpurun_out/launches.csv (ncu launch list) and the ncu --set full capture (gpurun_out/prof_<tag>.ncu-rep, or its
`--page raw --csv` export gpurun_out/prof_<tag>_raw.csv made on the GPU box) into the tracked summaries under
profiles/: <tag>_launches_summary.txt, <tag>_ncu_full_summary.csv, <tag>_ncu_traffic.json (DRAM bytes per launch of the
largest launch of each kernel: bench.py's `traffic`).  Usage: python scripts/summarize_profiles.py r02"""
import collections, csv, io, json, os, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_dir = os.path.join(ROOT, "profiles")
# ---- launch list
lines = [l for l in open(os.path.join(ROOT, "gpurun_out", "launches.csv")) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    try:
        name, v, unit = row["Kernel Name"], float(row["Metric Value"].replace(",", "")), row.get("Metric Unit", "")
    except Exception:
        continue
    v = v / 1e6 if unit == "ns" else v / 1e3 if unit in ("us", "usecond") else v
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(os.path.join(out_dir, "%s_launches_summary.txt" % tag), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py --steps 3 --warmup 3 --no-extras\n")
    f.write("# (cold-cache, serialised launches: compare SHARES with bench.py's `kernels`, not absolutes)\n")
    f.write("%-90s %5s %10s %9s %6s\n" % ("kernel", "n", "total_ms", "avg_ms", "share"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%-90s %5d %10.3f %9.4f %5.1f%%\n" % (k[:90], a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
    f.write("total %.3f ms over %d launches\n" % (tot, sum(a[0] for a in agg.values())))
# ---- full capture
rep = os.path.join(ROOT, "gpurun_out", "prof_%s.ncu-rep" % tag)
rawcsv = os.path.join(ROOT, "gpurun_out", "prof_%s_raw.csv" % tag)
if os.path.exists(rawcsv) or os.path.exists(rep):
    raw = open(rawcsv).read() if os.path.exists(rawcsv) else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
            "sm__cycles_elapsed.avg.per_second", "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
    idx = []
    for w in want:                                  # the plain metric if the capture has it, else a section-prefixed one
        hit = [i for i, h in enumerate(hdr) if h == w] or [i for i, h in enumerate(hdr) if h.endswith("." + w)]
        if hit:
            idx.append(hit[0])
    with open(os.path.join(out_dir, "%s_ncu_full_summary.csv" % tag), "w") as f:
        wr = csv.writer(f)
        wr.writerow([hdr[i] for i in idx]); wr.writerow([units[i] for i in idx])
        for r in rows[2:]:
            wr.writerow([r[i] for i in idx])
    # DRAM traffic per launch (GB) of the largest launch of each kernel, keyed as bench.py's roofline entries
    names = {"conv_pool_tc_kernel": "conv_pool_tc", "dw2_tc_kernel": "dw2_tc", "dx2_tc_kernel": "dx2_tc", "adam_table_kernel": "adam_table",
             "emb_gather_bf16_drop_kernel": "emb_gather", "seg_reduce_tree_kernel": "emb_scatter", "gemm_tf32_async_kernel": "dense",
             "defer_adam_rows_kernel": "adam_defer_rows", "defer_settle_list_kernel": "adam_defer_settle_tokens"}
    ki, ri, wi, ti = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    def gb(v, u):
        v = float(v.replace(",", ""))
        return v * {"Gbyte": 1.0, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(u, 1.0)
    traffic = {}
    for r in rows[2:]:
        for frag, key in names.items():
            if frag in r[ki]:
                t = gb(r[ri], units[ri]) + gb(r[wi], units[wi])
                if t > traffic.get(key, 0.0):
                    traffic[key] = round(t, 4)
    with open(os.path.join(out_dir, "%s_ncu_traffic.json" % tag), "w") as f:
        json.dump(traffic, f, indent=1, sort_keys=True)
print("ok")
