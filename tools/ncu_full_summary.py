"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one row per profiled launch of an `ncu --set full` capture) into a
markdown table: one line per launch, in launch order, with the counters DESIGN.md and VERDICT.md argue with.

usage: python tools/ncu_full_summary.py gpurun_out/r2_full_final_raw.csv > profiles/r2_ncu_full_final.md
"""
import csv
import re
import sys

STALLS = ["barrier", "branch_resolving", "dispatch_stall", "drain", "lg_throttle", "long_scoreboard", "math_pipe_throttle", "membar",
          "mio_throttle", "misc", "no_instruction", "not_selected", "short_scoreboard", "sleeping", "tex_throttle", "wait"]


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return float("nan")


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def get(r, name):
        return num(r[col[name]]) if name in col else float("nan")

    def unit(name):
        return units[col[name]] if name in col else ""

    print(f"# ncu --set full, one row per launch ({len(data)} launches, launch order)\n")
    print("Cold-cache, serialised replays (`--clock-control none`): durations are NOT the in-graph times; the counters are what this file is for.\n")
    print("| # | kernel | grid x block | us | regs | waves | occ % | issue % | M inst | dram % | dram MB r+w | L2 hit % | tensor % | top stalls (warps per issue) |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    tot_us = tot_inst = 0.0
    for i, r in enumerate(data):
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).split("::")[-1]
        dur = get(r, "gpu__time_duration.sum")
        dur *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit("gpu__time_duration.sum"), 1e-3)
        inst = get(r, "smsp__inst_executed.sum")
        rd, wr = get(r, "dram__bytes_read.sum"), get(r, "dram__bytes_write.sum")
        scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
        mb = rd * scale.get(unit("dram__bytes_read.sum"), 1e-6) + wr * scale.get(unit("dram__bytes_write.sum"), 1e-6)
        st = sorted(((get(r, f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"), s) for s in STALLS), reverse=True)
        st = ", ".join(f"{s} {v:.1f}" for v, s in st[:3] if v == v)
        tensor = get(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
        print(f"| {i} | {name[:44]} | {r[col['launch__grid_size']]} x {r[col['launch__block_size']]} | {dur:.1f} | "
              f"{int(get(r, 'launch__registers_per_thread'))} | {get(r, 'launch__waves_per_multiprocessor'):.2f} | "
              f"{get(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f} | {get(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.0f} | "
              f"{inst / 1e6:.2f} | {get(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | {mb:.2f} | "
              f"{get(r, 'lts__t_sector_hit_rate.pct'):.0f} | {tensor:.0f} | {st} |")
        tot_us += dur
        tot_inst += inst
    print(f"\nTotal: {tot_us:.1f} us serialised, {tot_inst / 1e6:.1f} M warp instructions "
          f"({tot_inst / (148 * 4 * 1.965e3):.1f} us at the chip's peak issue rate, 148 SMs x 4 schedulers x 1.965 GHz).")


if __name__ == "__main__":
    main()
