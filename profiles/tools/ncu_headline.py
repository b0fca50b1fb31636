"""Prints the headline counters of one kernel from `ncu -i <rep> --page raw --csv` output (profiles/README.md table).
usage: python profiles/tools/ncu_headline.py profiles/r1_ncu_raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, unit, val = rows[0], rows[1], rows[-1]
m = {h: (val[i], unit[i]) for i, h in enumerate(hdr)}


def g(name):
    v, u = m.get(name, ("", ""))
    return v.replace(",", ""), u


for name in ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
             "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
             "sm__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
             "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
             "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__icc_request_hit_rate.pct",
             "sm__icc_requests.sum", "gcc__cache_requests_type_instruction.sum",
             "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed",
             "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
             "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
             "l1tex__t_sector_hit_rate.pct"]:
    v, u = g(name)
    if v:
        print("%-75s %s %s" % (name, v, u))
stalls = []
for h in hdr:
    if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
        try:
            stalls.append((float(g(h)[0]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        except ValueError:
            pass
print("stalls (warps per issue):", ", ".join("%s %.1f" % (n, v) for v, n in sorted(stalls, reverse=True)[:6]))
