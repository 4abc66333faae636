"""Summarises an .ncu-rep (ncu --set full) into text: per-launch headline metrics + the SASS lines with most stall samples.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/xxx.txt"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
traffic_out = sys.argv[sys.argv.index("--traffic") + 1] if "--traffic" in sys.argv else None   # merge DRAM bytes per launch into this json
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.max",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
names = []
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    names.append(name)
    print("==", name[:110])
    for w in WANT:
        if w in idx:
            print(f"   {w:68s} {r[idx[w]]:>16s} {units[idx[w]]}")
if traffic_out:
    import json, os, re
    UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def bench_name(n):   # the names bench.py's `kernels` entries carry
        if "contract_x3_pair_kernel" in n:
            return "contract_x3_pair_kernel"
        if "homotopy_algebra_kernel" in n:
            return "homotopy_algebra_kernel"
        m = re.search(r"homotopy_scan_kernel<([^>]*)>", n)   # <T, THREADS, NCH, GRAMQ, EXACT>
        if m:
            par = [p.replace("(bool)", "").strip() for p in m.group(1).split(",")]
            return "homotopy_scan_kernel (q from Gram rows)" if par[3] in ("1", "true") else "homotopy_scan_kernel (Q streamed)"
        return None
    cfg_name = sys.argv[sys.argv.index("--config") + 1] if "--config" in sys.argv else "C2"
    whole = json.load(open(traffic_out)) if os.path.exists(traffic_out) else {}
    out = whole.setdefault(cfg_name, {"bytes_per_launch": {}, "source": {}})
    for r in rows[2:]:
        bn = bench_name(r[idx["Kernel Name"]])
        if bn is None:
            continue
        tot = sum(float(r[idx[k]].replace(",", "")) * UNIT[units[idx[k]]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        grid = r[idx["launch__grid_size"]]
        if bn == "homotopy_algebra_kernel" and tot < out["bytes_per_launch"].get(bn, 0):
            continue      # the late-iteration launch (the one that gathers p = A_G d) is the figure; a Gram-iteration launch moves 1/8 of it
        out["bytes_per_launch"][bn] = tot
        out["source"][bn] = f"{os.path.basename(rep)}: {r[idx['Kernel Name']][:70]}, grid {grid}, {r[idx['gpu__time_duration.sum']]} {units[idx['gpu__time_duration.sum']]}"
    json.dump(whole, open(traffic_out, "w"), indent=1)
for pat in sorted(set(n.split("(")[0].split("<")[0].split("::")[-1].split()[-1] for n in names)):
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = next((i for i, r in enumerate(rows) if r and r[0] == "Address"), None)
    if h is None:
        continue
    hdr = rows[h]
    si = hdr.index("Warp Stall Sampling (All Samples)")
    stall_cols = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
    data, tot_by = [], {hdr[i]: 0 for i in stall_cols}
    seen = set()
    for r in rows[h + 1:]:
        if len(r) <= si or r[0] in seen or r[0] == "Address":
            continue
        seen.add(r[0])
        try:
            n = int(r[si])
        except ValueError:
            continue
        data.append((n, r[1].strip(), r))
        for i in stall_cols:
            try:
                tot_by[hdr[i]] += int(r[i])
            except ValueError:
                pass
    tot = sum(d[0] for d in data) or 1
    print(f"\n== stall samples, kernel ~{pat}: total {tot} (first launch)")
    print("   by reason:", ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in sorted(tot_by.items(), key=lambda kv: -kv[1])[:8]))
    for n, sass, r in sorted(data, key=lambda d: -d[0])[:int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 14]:
        print(f"   {100 * n / tot:5.1f}%  {sass[:100]}")
