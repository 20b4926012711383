"""Summarise an `ncu --set full` report as a markdown table (one column per captured launch).

    ncu -i gpurun_out/prof_x.ncu-rep --page raw --csv > /tmp/x.csv
    python tools/ncu_summary.py /tmp/x.csv [--title "..."] >> profiles/r2/ncu_x.md

Only the metrics the profiling recipe asks for are kept (duration, tensor-pipe activity, DRAM / L2 traffic and throughput,
occupancy, registers, shared memory, issue-slot utilisation, top stall reasons)."""
import csv
import sys

KEEP = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem)"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs)"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active % (of active cycles)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (of elapsed)"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor instructions"),
    ("sm__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("smsp__inst_executed.sum", "warp instructions executed"),
    ("sm__inst_executed_pipe_xu.sum", "XU (MUFU) instructions"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM bytes read"),
    ("dram__bytes_write.sum", "DRAM bytes written"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1/TEX throughput %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared-memory wavefronts"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall: long scoreboard"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall: long scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall: barrier / issue"),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall: membar / issue"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall: short scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall: wait / issue"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall: math pipe throttle / issue"),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "stall: MIO throttle / issue"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall: LG throttle / issue"),
    ("smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio", "stall: sleeping / issue"),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "stall: no instruction / issue"),
    ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stall: branch resolving / issue"),
    ("smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "stall: dispatch / issue"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall: not selected / issue"),
    ("smsp__average_warps_issue_stalled_selected_per_issue_active.ratio", "selected / issue"),
]


def main() -> None:
    path = sys.argv[1]
    title = sys.argv[sys.argv.index("--title") + 1] if "--title" in sys.argv else path
    rows = list(csv.reader(open(path, newline="")))
    # the raw page starts with optional "==PROF==" lines; the header row contains "Kernel Name"
    hdr_i = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[hdr_i], rows[hdr_i + 1]
    data = [r for r in rows[hdr_i + 2:] if len(r) == len(hdr)]
    col = {name: i for i, name in enumerate(hdr)}
    print(f"## {title}\n")
    names = [r[col["Kernel Name"]][:70] for r in data]
    print("| metric | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |")
    print("|---|" + "---:|" * len(data))
    print("| kernel | " + " | ".join(f"`{n}`" for n in names) + " |")
    for key, label in KEEP:
        if key not in col:
            continue
        u = units[col[key]]
        vals = [r[col[key]] for r in data]
        print(f"| {label}" + (f" [{u}]" if u else "") + " | " + " | ".join(vals) + " |")
    print()


if __name__ == "__main__":
    main()
