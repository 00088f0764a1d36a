"""Print the handful of ncu metrics this project reads from a .ncu-rep (usage: ncu_summary.py file.ncu-rep)."""
import csv
import subprocess
import sys

WANT = [
    "Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_pipe_lsu.sum",
    "sm__inst_executed_pipe_tensor.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
]
STALL = "smsp__average_warps_issue_stalled_"


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units = rows[0], rows[1]
    for row in rows[2:]:
        print("----")
        for w in WANT:
            if w in head:
                i = head.index(w)
                print("%-70s %s %s" % (w, row[i][:90], units[i]))
        stalls = [(float(row[i]), head[i][len(STALL):].replace("_per_issue_active.ratio", ""))
                  for i in range(len(head)) if head[i].startswith(STALL) and head[i].endswith("_per_issue_active.ratio")
                  and row[i] not in ("", "n/a")]
        for v, n in sorted(stalls, reverse=True)[:7]:
            print("  stall %-28s %.3f" % (n, v))


if __name__ == "__main__":
    main(sys.argv[1])
