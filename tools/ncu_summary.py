"""Summarise an .ncu-rep (read here, no GPU needed) into a small text table for profiles/.
Usage: python tools/ncu_summary.py gpurun_out/X.ncu-rep > profiles/X.summary.txt"""
import csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__registers_per_thread", "regs/thread"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs), blocks/SM"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "avg active threads / warp instruction (warp execution efficiency x32)"),
    ("smsp__thread_inst_executed_pred_on_per_inst_executed.ratio", "avg not-predicated-off threads / warp instruction"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slot utilisation %"),
    ("sm__inst_executed.avg.per_cycle_elapsed", "warp instructions / cycle / SM"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 sector hit %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1/TEX throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 sector hit %"),
    ("lts__t_sectors.sum", "L2 sectors (32 B)"),
    ("l1tex__t_sectors.sum", "L1 sectors (32 B)"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("smsp__sass_average_branch_targets_threads_uniform.pct", "branch uniformity %"),
]
STALLS = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    print(f"== {name}  (launch id {r[col['ID']]})")
    dur = None
    for k, label in KEYS:
        if k in col:
            v, u = r[col[k]], units[col[k]]
            print(f"  {label:75s} {v} {u}")
            if k == "gpu__time_duration.sum":
                dur = float(v.replace(",", "")) * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(u, 1.0)

    def num(k):
        if k not in col:
            hits = [h for h in col if h.endswith(k)]
            if not hits: return float('nan')  # section-limited capture: metric absent
            k = hits[0]
        v, u = r[col[k]].replace(",", ""), units[col[k]]
        return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1.0)
    if dur:
        print(f"  {'achieved L1 GB/s (l1tex__t_sectors x 32 B / duration)':75s} {num('l1tex__t_sectors.sum') * 32 / dur / 1e9:.1f}")
        print(f"  {'achieved L2 GB/s (lts__t_sectors x 32 B / duration)':75s} {num('lts__t_sectors.sum') * 32 / dur / 1e9:.1f}")
        print(f"  {'achieved HBM GB/s ((dram read + write) / duration)':75s} {(num('dram__bytes_read.sum') + num('dram__bytes_write.sum')) / dur / 1e9:.1f}")
        print(f"  {'DRAM traffic per launch, bytes':75s} {num('dram__bytes_read.sum') + num('dram__bytes_write.sum'):.4g}")
    st = sorted(((float(r[col[h]]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]) for h in STALLS if r[col[h]]), reverse=True)
    print("  warp stall reasons (avg warps stalled per issue-active cycle): " + ", ".join(f"{n} {v:.2f}" for v, n in st[:7]))
