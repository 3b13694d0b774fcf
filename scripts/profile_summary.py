"""Turn the ncu captures brought back in gpurun_out/ into the tracked summaries under profiles/.

    python scripts/profile_summary.py gpurun_out/r1_full.ncu-rep gpurun_out/r1_launches.csv r1

writes profiles/<tag>_ncu_full_summary.txt, profiles/<tag>_launch_list_summary.txt and
profiles/ncu_summary.json (bench.py reads `onesweep_dram_bytes_per_launch` from the latter for
roofline.traffic).  Dev tool: runs in the build container (ncu -i needs no GPU)."""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_st.sum", "lts__t_sector_hit_rate.pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
]


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units = rows[0], rows[1]
    return head, units, rows[2:]


def to_float(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    rep, launches, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    head, units, rows = raw_rows(rep)
    col = {h: i for i, h in enumerate(head)}
    picked = {}
    for r in rows:
        name = r[col["Kernel Name"]]
        kind = "onesweep" if "onesweep" in name else ("histogram" if "histogram" in name else None)
        if kind:
            picked[kind] = r  # the last capture of each kind (warm)
    lines, summary = [], {"round": tag, "command": "ncu --set full --import-source on --clock-control none -k regex:'onesweep|histogram' -s 5 -c 5 scripts/devbench 28 1 uniform 1 0",
                          "workload": "C2: 2^28 u32 key + u32 value pairs", "raw": {}}
    for kind, r in picked.items():
        lines.append(f"== {kind} ({r[col['Kernel Name']][:110]}) ==")
        summary["raw"][kind] = {}
        for k in KEEP:
            if k in col:
                lines.append(f"  {k:86s} {units[col[k]]:10s} {r[col[k]]}")
                summary["raw"][kind][k] = r[col[k]]
        stalls = [(h, to_float(r[i])) for h, i in col.items() if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
        stalls = [(h, v) for h, v in stalls if v]
        tot = sum(v for _, v in stalls) or 1.0
        top = sorted(stalls, key=lambda t: -t[1])[:8]
        lines.append("  warp-stall samples: " + ", ".join(f"{h.rsplit('stalled_', 1)[1]} {100 * v / tot:.1f}%" for h, v in top))

        def scaled(key):
            v, u = to_float(r[col[key]]), units[col[key]].lower()
            return v * {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
        rd, wr = scaled("dram__bytes_read.sum"), scaled("dram__bytes_write.sum")
        summary[f"{kind}_dram_read_bytes"] = rd
        summary[f"{kind}_dram_write_bytes"] = wr
        summary[f"{kind}_dram_bytes_per_launch"] = rd + wr
    n = 1 << 28
    summary["onesweep_algorithmic_bytes_per_launch"] = 2 * n * 8
    summary["histogram_algorithmic_bytes"] = n * 4
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_full_summary.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    with open(os.path.join(ROOT, "profiles", "ncu_summary.json"), "w") as f:
        json.dump(summary, f, indent=1)

    # launch list of the bench command: per-kernel count / total / share
    rows = [r for r in csv.reader(open(launches)) if len(r) > 10]
    head = rows[0]
    col = {h: i for i, h in enumerate(head)}
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if r[col["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v, u = to_float(r[col["Metric Value"]]), r[col["Metric Unit"]]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
        a = agg.setdefault(r[col["Kernel Name"]][:64], [0, 0.0])
        a[0] += 1
        a[1] += ms
    tot = sum(a[1] for a in agg.values())
    out = ["ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-log2n 22",
           "(first 400 launches; cold-cache serialised times: compare SHARES)",
           f"{'kernel':64s} {'count':>6s} {'total ms':>10s} {'share':>7s}"]
    for k, a in sorted(agg.items(), key=lambda t: -t[1][1]):
        out.append(f"{k:64s} {a[0]:6d} {a[1]:10.3f} {100 * a[1] / tot:6.1f}%")
    thrs = sum(a[1] for k, a in agg.items() if "thrs_b200" in k)
    one = sum(a[1] for k, a in agg.items() if "onesweep" in k)
    out.append(f"onesweep share of the library's own kernels: {100 * one / thrs:.1f}% (the rest: histogram, probe)")
    with open(os.path.join(ROOT, "profiles", f"{tag}_launch_list_summary.txt"), "w") as f:
        f.write("\n".join(out) + "\n")
    print("\n".join(lines))
    print("\n".join(out[:12]))


if __name__ == "__main__":
    main()
