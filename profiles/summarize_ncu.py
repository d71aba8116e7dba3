"""Turn `ncu -i X.ncu-rep --page raw --csv` output into a compact metric,unit,value-per-launch table.
usage: python profiles/summarize_ncu.py raw.csv out.csv [substring-filter ...]"""
import csv
import sys

DEFAULT = ("Kernel Name", "gpu__time_duration", "dram__bytes", "dram__throughput", "lts__t_bytes.sum",
           "lts__t_sector_hit_rate", "sm__throughput", "smsp__inst_executed.sum", "smsp__issue_active",
           "sm__warps_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared", "l1tex__data_bank_conflicts",
           "l1tex__throughput", "launch__", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
           "issue_stalled", "sm__pipe_", "sm__inst_executed_pipe", "smsp__thread_inst_executed_per_inst",
           "smsp__sass_average_branch", "sm__sass_inst_executed_op_shared", "sm__sass_inst_executed_op_global",
           "l1tex__t_sector_hit_rate", "l1tex__t_bytes")


def main():
    src, dst = sys.argv[1], sys.argv[2]
    filt = tuple(sys.argv[3:]) or DEFAULT
    rows = list(csv.reader(open(src)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
        for i, k in enumerate(hdr):
            if any(s in k for s in filt):
                w.writerow([k, units[i]] + [r[i] for r in data])


if __name__ == "__main__":
    main()
