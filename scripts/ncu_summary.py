#!/usr/bin/env python
"""Summarise one `ncu --set full` capture (gpurun_out/*.ncu-rep) into a small JSON/markdown pair
under profiles/: duration, DRAM bytes, throughput percentages, occupancy, instruction mix.

    python scripts/ncu_summary.py gpurun_out/prof_frame.ncu-rep profiles/r01_frame_kernel_<tag>
"""
import collections
import csv
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    kernels = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                d[k] = (r[i], units[i])
        kernels.append(d)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(src.splitlines()))
    mix = collections.Counter()
    if len(srows) > 2:
        h = srows[1]
        if "Instructions Executed" in h:
            ie, isrc = h.index("Instructions Executed"), h.index("Source")
            for r in srows[2:]:
                if len(r) <= ie or not r[ie].isdigit() or not r[isrc].split():
                    continue  # a second kernel's header rows in a multi-kernel report
                toks = r[isrc].split()
                op = toks[1] if toks[0].startswith("@") else toks[0]
                mix[op.split(".")[0]] += int(r[ie])
    st_cols = [(i, k.split("issue_stalled_")[1].split("_per_issue")[0]) for i, k in enumerate(hdr)
               if "issue_stalled" in k and "per_issue_active" in k]
    for d, r in zip(kernels, rows[2:]):
        d["stalls_per_issue"] = dict(sorted(((n, round(float(r[i]), 2)) for i, n in st_cols if r[i]), key=lambda t: -t[1])[:8])
    k0 = kernels[0]
    dur_ms = float(k0["gpu__time_duration.sum"][0]) * {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}[k0["gpu__time_duration.sum"][1]]

    def gb(key):
        v, u = k0[key]
        return float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}[u]
    traffic = gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum")
    summary = {"report": rep, "kernel": k0["kernel"], "duration_ms_under_ncu": dur_ms, "dram_bytes_per_launch": traffic,
               "dram_read_bytes": gb("dram__bytes_read.sum"), "dram_write_bytes": gb("dram__bytes_write.sum"),
               "metrics": {k: v for k, v in k0.items() if k not in ("kernel", "stalls_per_issue")},
               "stalls_per_issue": k0["stalls_per_issue"],
               "other_kernels_in_report": [{"kernel": k["kernel"], "duration": k.get("gpu__time_duration.sum"),
                                            "dram_read": k.get("dram__bytes_read.sum"), "dram_write": k.get("dram__bytes_write.sum"),
                                            "warp_instructions": k.get("smsp__inst_executed.sum"),
                                            "issue_active_pct": k.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                                            "stalls_per_issue": k["stalls_per_issue"]} for k in kernels[1:]],
               "warp_instructions_by_opcode": dict(mix.most_common(25)), "warp_instructions_total": sum(mix.values())}
    json.dump(summary, open(out + ".json", "w"), indent=1)
    with open(out + ".md", "w") as f:
        f.write(f"# ncu --set full: {k0['kernel']}\n\nreport: `{rep}` (scratch, not committed)\n\n")
        f.write(f"- duration under ncu: {dur_ms:.4f} ms\n- DRAM traffic: {traffic / 1e9:.3f} GB per launch "
                f"(read {gb('dram__bytes_read.sum') / 1e9:.3f}, write {gb('dram__bytes_write.sum') / 1e9:.3f})\n")
        for k, (v, u) in summary["metrics"].items():
            f.write(f"- {k}: {v} {u}\n")
        f.write(f"- stalls per issue: {summary['stalls_per_issue']}\n")
        for k in summary["other_kernels_in_report"]:
            f.write(f"\n## also in the report: {k['kernel']}\n\n- duration {k['duration']}, DRAM read {k['dram_read']}, write {k['dram_write']}, "
                    f"warp instructions {k['warp_instructions']}, issue active {k['issue_active_pct']}\n- stalls per issue: {k['stalls_per_issue']}\n")
        f.write("\n## warp instructions executed by opcode (all kernels of the report)\n\n")
        tot = max(sum(mix.values()), 1)
        for op, n in mix.most_common(25):
            f.write(f"- {op}: {n} ({100.0 * n / tot:.1f}%)\n")
    print(json.dumps({k: summary[k] for k in ("kernel", "duration_ms_under_ncu", "dram_bytes_per_launch", "warp_instructions_total")}))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
