"""profiles/roofs_<workload>_f<bits>.json from an `ncu --set full` summary (tools/ncu_summary.py output): the three roofs bench.py
reports next to the HBM fraction — instruction issue, L1 / shared-memory data pipe, DRAM — for the dominant kernel.
Usage: python tools/ncu_roofs.py SUMMARY.json KERNEL_SUBSTRING KERNEL_CLASS OUT.json"""

import json
import sys


def main():
    summary, needle, kernel_class, out = sys.argv[1:5]
    d = json.load(open(summary))
    rows = [r for r in d["launches"] if needle in r["kernel"]]
    if not rows:
        raise SystemExit(f"no launch of {needle!r} in {summary}")
    r = max(rows, key=lambda x: x.get("gpu__time_duration.sum", 0.0))  # the full-size launch (not a pipelined chunk)
    rec = {
        "kernel": r["kernel"][:120], "kernel_class": kernel_class, "source": summary,
        "duration_us_under_ncu": r.get("gpu__time_duration.sum"),
        "dram_bytes_per_launch": r.get("dram_bytes_total"),
        "dram_pct": r.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "issue_active_pct": r.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "l1_data_pipe_pct": r.get("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "fma_pipe_pct": r.get("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        "alu_pipe_pct": r.get("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "xu_pipe_pct": r.get("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        "fma_pipe_cycles_pct": r.get("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        "xu_pipe_cycles_pct": r.get("sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active"),
        "alu_pipe_cycles_pct": r.get("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
        "warps_active_pct": r.get("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "warp_instructions": r.get("smsp__inst_executed.sum"),
        "threads_per_instruction": r.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "registers_per_thread": r.get("launch__registers_per_thread"),
    }
    json.dump(rec, open(out, "w"), indent=1)
    print(json.dumps(rec, indent=1))


if __name__ == "__main__":
    main()
