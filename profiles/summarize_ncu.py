"""Reduce an `ncu --set full` report (read with `ncu -i X.ncu-rep --page raw --csv`) to the
handful of metrics the roofline discussion in DESIGN.md uses, and the opcode mix of the
source page.  Usage: python profiles/summarize_ncu.py gpurun_out/X.ncu-rep profiles/X_summary.json"""
import collections
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__sass_inst_executed_op_tmem_ldt.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main(rep, dst):
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    launches = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")][:120]}
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                try:
                    d[k] = float(r[i].replace(",", ""))
                except ValueError:
                    d[k] = r[i]
                d[k + "__unit"] = units[i]
        launches.append(d)
    src = ncu_csv(rep, "source", ("--print-source", "sass"))
    ops = collections.Counter()
    total = 0
    if len(src) > 2:
        h = src[1]
        si, ei = h.index("Source"), h.index("Instructions Executed")
        for r in src[2:]:
            if r and r[0] == "Kernel Name":
                break
            if len(r) <= ei:
                continue
            toks = [t for t in r[si].split() if not t.startswith("@")]
            n = int(r[ei])
            ops[toks[0].split(".")[0]] += n
            total += n
    json.dump({"report": rep, "launches": launches, "warp_instructions_first_launch": total,
               "opcode_mix_first_launch": dict(ops.most_common(24))}, open(dst, "w"), indent=1)
    print("wrote", dst)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
