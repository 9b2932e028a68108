"""Summarise an .ncu-rep for profiles/: a fixed set of raw metrics per launch (JSON) and, with --source, the
warp-stall samples and executed warp instructions by opcode of the first launch (text).

    python scratch/ncu_summary.py gpurun_out/x.ncu-rep profiles/x_summary.json [--source profiles/x_source.txt] [--note "..."]
"""
import csv, io, json, subprocess, sys, collections, re

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active", "smsp__inst_executed.sum",
    "sm__cycles_active.avg", "sm__cycles_active.max", "sm__cycles_elapsed.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.avg.pct_of_peak_sustained_elapsed",
    "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_not_selected",
    "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_wait",
    "smsp__pcsamp_warps_issue_stalled_dispatch_stall", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
    "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_branch_resolving",
    "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_no_instructions",
]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    note = sys.argv[sys.argv.index("--note") + 1] if "--note" in sys.argv else ""
    rows = page(rep, "raw")
    hdr, units = rows[0], rows[1]
    launches = []
    for vals in rows[2:]:
        d = {"Kernel Name": vals[hdr.index("Kernel Name")]}
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                d[m] = f"{vals[i]} {units[i]}".strip()
        launches.append(d)
    json.dump({"report": rep.split("/")[-1], "note": note, "launches": launches}, open(dst, "w"), indent=1)
    if "--source" in sys.argv:
        out = sys.argv[sys.argv.index("--source") + 1]
        src = page(rep, "source")
        by = collections.defaultdict(lambda: [0, 0])
        seen, hot = set(), []
        col, first_kernel = None, None
        for r in src:
            if r and r[0] == "Kernel Name":
                if first_kernel is None:
                    first_kernel = r[1]
                elif r[1] != first_kernel or col is not None:
                    break          # next launch (or the second view of the same one)
                continue
            if r and r[0] == "Address":
                col = {name: i for i, name in enumerate(r)}
                continue
            if col is None or len(r) < len(col):
                continue
            key = (r[col["Address"]], r[col["Source"]])
            if key in seen:       # count each SASS address once
                continue
            seen.add(key)
            text = r[col["Source"]].strip()
            op = re.sub(r"^@!?U?P\d+\s+", "", text).split(" ")[0].split(".")[0]
            try:
                s_, e_ = int(r[col["# Samples"]] or 0), int(r[col["Instructions Executed"]] or 0)
            except ValueError:
                continue
            by[op][0] += s_
            by[op][1] += e_
            hot.append((s_, text))
        tot = sum(v[0] for v in by.values()) or 1
        with open(out, "w") as f:
            f.write(f"ncu --page source of {rep.split('/')[-1]} (first launch): warp-stall samples and executed warp instructions by opcode, each SASS address counted once\n")
            if note:
                f.write(note + "\n")
            f.write(f"total samples {tot}\nopcode  stall_samples  share  executed_warp_instr\n")
            for op, (s_, e_) in sorted(by.items(), key=lambda kv: -kv[1][0])[:18]:
                f.write(f"{op:12s} {s_:10d} {100 * s_ / tot:5.1f}% {e_:14d}\n")
            f.write("hottest instructions:\n")
            for s_, text in sorted(hot, reverse=True)[:12]:
                f.write(f"{s_:8d} {100 * s_ / tot:4.1f}%  {text}\n")


if __name__ == "__main__":
    main()
