"""Summarise an ncu report (read here, no GPU needed): python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/out.md"""
import csv, io, subprocess, sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 % of peak"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/TEX % of peak"),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1 data-pipe wavefronts % of peak"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM % of peak"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / warp instruction"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"),
]
STALLS = ["long_scoreboard", "wait", "not_selected", "short_scoreboard", "branch_resolving", "math_pipe_throttle",
          "lg_throttle", "no_instructions", "barrier", "mio_throttle", "dispatch_stall", "drain", "imc_miss", "membar", "sleeping", "tex_throttle", "selected"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = [f"# ncu --set full summary of `{rep}`", "",
             "Per-launch values are cold-cache and serialised by the profiler (compare shares, not absolutes).", ""]
    for r in rows[2:]:
        lines.append(f"## {r[idx['Kernel Name']]}  (launch id {r[idx['ID']]})")
        lines.append("")
        lines.append("| metric | value |")
        lines.append("|---|---|")
        for k, name in KEYS:
            if k in idx:
                lines.append(f"| {name} (`{k}`) | {r[idx[k]]} {units[idx[k]]} |")
        tot = 0.0; st = []
        for s in STALLS:
            k = f"smsp__pcsamp_warps_issue_stalled_{s}"
            if k in idx and r[idx[k]]:
                v = float(r[idx[k]]); st.append((v, s)); tot += v
        if tot > 0:
            top = ", ".join(f"{s} {100 * v / tot:.0f}%" for v, s in sorted(st, reverse=True)[:5])
            lines.append(f"| warp-state samples (top 5) | {top} |")
        lines.append("")
    open(out, "w").write("\n".join(lines))
    print(f"wrote {out}: {len(rows) - 2} launches")


if __name__ == "__main__":
    main()
