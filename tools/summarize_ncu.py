"""Summarise ncu outputs brought back in gpurun_out/ into small tracked files under profiles/.

    python tools/summarize_ncu.py r1          # reads gpurun_out/r1_launches_*.csv and gpurun_out/r1_full_*.ncu-rep
"""
import collections, csv, glob, json, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
out_dir = os.path.join(ROOT, "profiles")
os.makedirs(out_dir, exist_ok=True)


def short(name):
    name = re.sub(r"^void ", "", name)
    name = name.replace("<unnamed>::", "")
    name = re.sub(r"\((bool|int|unsigned int)\)", "", name)          # template arguments print as (bool)0, (int)256
    return name.split("(")[0][:90]


for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_launches_*.csv"))):
    wl = os.path.basename(path)[len(tag) + 10:-4]
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            agg.setdefault(short(r[ki]), []).append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
    tot = sum(sum(v) for v in agg.values())
    lines = [f"# ncu launch list, workload {wl}: python bench.py --workload {wl} --steps 2 --warmup 3 --no-cpu-baseline --no-stages --no-workloads",
             "# ncu --metrics gpu__time_duration.sum --clock-control none (per-launch times are cold-cache and serialised: compare SHARES)",
             f"{'kernel':<92s} {'launches':>8s} {'mean_ns':>10s} {'share':>7s}"]
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{k:<92s} {len(v):8d} {sum(v) / len(v):10.0f} {sum(v) / tot:7.3f}")
    open(os.path.join(out_dir, f"{tag}_launches_{wl}.txt"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__icc_request_hit_rate.pct", "sm__cycles_elapsed.max"]
traffic = {}
for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_full_*.ncu-rep"))):
    wl = os.path.basename(path)[len(tag) + 6:-8]
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ni = hdr.index("Kernel Name")
    res = {"workload": wl, "command": f"ncu --set full --clock-control none --import-source on (bench.py --workload {wl} --steps 2 --warmup 3)",
           "launches": []}
    for r in rows[2:]:
        ent = {"kernel": short(r[ni])}
        for i, h in enumerate(hdr):
            if h in WANT:
                ent[h] = f"{r[i]} {units[i]}".strip()
        stalls = {}
        for i, h in enumerate(hdr):
            if "pcsamp_warps_issue_stalled" in h and "not_issued" not in h:
                try:
                    stalls[h.replace("smsp__pcsamp_warps_issue_stalled_", "")] = float(r[i])
                except ValueError:
                    pass
        s = sum(stalls.values()) or 1.0
        ent["stall_share_top"] = {k: round(v / s, 3) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:6]}
        res["launches"].append(ent)

        def tobytes(val, unit):
            mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            return float(val) * mul
        try:
            rd = tobytes(r[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_read.sum")])
            wr = tobytes(r[hdr.index("dram__bytes_write.sum")], units[hdr.index("dram__bytes_write.sum")])
            traffic.setdefault(wl, {}).setdefault(ent["kernel"], []).append(rd + wr)
        except (ValueError, KeyError):
            pass
    json.dump(res, open(os.path.join(out_dir, f"{tag}_ncu_full_{wl}.json"), "w"), indent=1)
    print(json.dumps(res, indent=1)[:1500])
for wl, t in traffic.items():
    for k in list(t):
        t[k] = sum(t[k]) / len(t[k])          # dram__bytes_read.sum + dram__bytes_write.sum per launch
tpath = os.path.join(out_dir, "ncu_traffic.json")
merged = json.load(open(tpath)) if os.path.exists(tpath) else {}
for wl, t in traffic.items():
    merged.setdefault(wl, {}).update(t)
json.dump(merged, open(tpath, "w"), indent=1)
print(traffic)
