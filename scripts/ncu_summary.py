"""Text summaries of ncu captures for profiles/ (run here, no GPU needed).

    python scripts/ncu_summary.py raw  <rep.ncu-rep>   # per-launch metric table of a --set full capture
    python scripts/ncu_summary.py list <launches.csv>  # per-kernel totals / shares of a gpu__time_duration launch list
"""
import collections
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__thread_inst_executed_pred_on_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    names = [r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")[:14] for r in body]
    print(f"# {rep}: {len(body)} launches; warp execution efficiency = thread_inst_executed_per_inst_executed / 32")
    print(f"{'metric':72s} {'unit':10s} " + " ".join(f"{n:>14s}" for n in names))
    for w in WANT:
        if w not in hdr:
            continue
        i = hdr.index(w)
        print(f"{w:72s} {units[i][:10]:10s} " + " ".join(f"{r[i][:14]:>14s}" for r in body))
    print("# stall reasons, warps per issue-active cycle (smsp__average_warps_issue_stalled_*_per_issue_active)")
    for i, h in enumerate(hdr):
        if "smsp__average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
            vals = [float(r[i] or 0) for r in body]
            if max(vals) > 0.1:
                print(f"{h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):72s} {'':10s} " + " ".join(f"{v:14.2f}" for v in vals))


def launch_list(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    i_name, i_val = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[i_val].replace(",", ""))
        except ValueError:
            continue
        n = r[i_name].split("(")[0].replace("void ", "")
        agg[n][0] += 1
        agg[n][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot / 1e6:.2f} ms of kernel time (ncu: serialised, cold cache; compare shares)")
    print(f"{'kernel':28s} {'launches':>8s} {'total us':>12s} {'share':>7s} {'avg us':>9s}")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:28s} {c:8d} {t / 1e3:12.1f} {t / tot:7.3f} {t / c / 1e3:9.2f}")


if __name__ == "__main__":
    (raw if sys.argv[1] == "raw" else launch_list)(sys.argv[2])
