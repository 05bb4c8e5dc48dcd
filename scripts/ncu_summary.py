"""Key figures of one `ncu --set full` capture (.ncu-rep), as text for profiles/.

    python scripts/ncu_summary.py capture.ncu-rep
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
]


def main():
    raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    for v in rows[2:]:
        d = {n: (v[i], u[i]) for i, n in enumerate(h)}
        print(d["Kernel Name"][0])
        for k in KEYS:
            if k in d:
                print(f"  {k:75s} {d[k][0]:>16s} {d[k][1]}")
        stalls = []
        for k in d:
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
                try:
                    stalls.append((float(d[k][0]), k.split("issue_stalled_")[1].split("_per_issue")[0]))
                except ValueError:
                    pass
        print("  warp stall cycles per issued instruction: " + ", ".join(f"{n} {x:.2f}" for x, n in sorted(stalls, reverse=True)[:6]))
        print()


if __name__ == "__main__":
    main()
