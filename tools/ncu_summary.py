#!/usr/bin/env python
"""Condenses `ncu -i X.ncu-rep --page raw --csv` dumps into the markdown tables committed under profiles/.

    python tools/ncu_summary.py TITLE steps_per_launch:raw.csv [...] > profiles/rNN_ncu_summary.md

steps_per_launch = K_local * H of the captured launch (for the derived per-rollout-step figures)."""
import csv
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "sm__cycles_elapsed.max", "sm__cycles_active.avg",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def load(path):
    rows = list(csv.reader(open(path)))
    head, units, vals = rows[0], rows[1], rows[2]
    d = {}
    for h, u, v in zip(head, units, vals):
        d[h.split(".", 2)[-1] if h.count(".") >= 2 and h.split(".")[1] in ("TriageCompute",) else h] = (v, u)
        d[h] = (v, u)
    return d


def num(s):
    try:
        return float(s.replace(",", ""))
    except Exception:
        return None


def evidence(path):
    """The few numbers bench.py attaches to its roofline object."""
    d = load(path)
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd, wr = d["dram__bytes_read.sum"], d["dram__bytes_write.sum"]
    g = lambda k: num(d[k][0]) if k in d else None
    return {"dram_bytes_per_launch": num(rd[0]) * mult[rd[1]] + num(wr[0]) * mult[wr[1]],
            "gpu_time_us": g("gpu__time_duration.sum"),
            "fma_pipe_cycles_active_pct": g("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
            "xu_pipe_pct": g("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
            "alu_pipe_pct": g("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
            "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "registers_per_thread": g("launch__registers_per_thread"), "source": path.split("/")[-1]}


def main():
    if sys.argv[1] == "--json":  # --json out.json key:raw.csv [...]  (merges into out.json)
        import json
        import os
        out = sys.argv[2]
        table = json.load(open(out)) if os.path.exists(out) else {}
        for spec in sys.argv[3:]:
            key, path = spec.split(":", 1)
            table[key] = evidence(path)
        json.dump(table, open(out, "w"), indent=1, sort_keys=True)
        return
    print("# %s\n" % sys.argv[1])
    for spec in sys.argv[2:]:
        steps, path = spec.split(":", 1)
        steps = float(steps)
        d = load(path)
        print("## %s\n" % path.split("/")[-1])
        print("kernel: `%s`  grid %s block %s\n" % (d["Kernel Name"][0][:110], d["Grid Size"][0], d["Block Size"][0]))
        print("| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in d:
                print("| %s | %s | %s |" % (k, d[k][0], d[k][1]))
        ti = num(d.get("smsp__thread_inst_executed.sum", ("", ""))[0])
        wi = num(d.get("smsp__inst_executed.sum", ("", ""))[0])
        if ti:
            print("| derived: thread instructions per rollout-step | %.1f | inst |" % (ti / steps))
        if wi:
            print("| derived: warp instructions x 32 per rollout-step | %.1f | inst |" % (wi * 32 / steps))
        rd, wr = d.get("dram__bytes_read.sum"), d.get("dram__bytes_write.sum")
        if rd and wr:
            mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tb = num(rd[0]) * mult.get(rd[1], 1) + num(wr[0]) * mult.get(wr[1], 1)
            print("| derived: DRAM traffic per launch (read + write) | %.1f | MB |" % (tb / 1e6))
            print("| derived: DRAM bytes per rollout-step | %.3f | B |" % (tb / steps))
        print()


if __name__ == "__main__":
    main()
