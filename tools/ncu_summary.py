"""Key metrics of an .ncu-rep (raw page) + top stall lines of the source page. Usage: ncu_summary.py rep [n_src_lines]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
nsrc = int(sys.argv[2]) if len(sys.argv) > 2 else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:]:
    print("=" * 100)
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:70s} {r[i]:>20s} {units[i]}")
    stall = [(hdr[i], r[i]) for i in range(len(hdr)) if "smsp__average_warp" in hdr[i] and "issue_stalled" in hdr[i] and hdr[i].endswith("_per_warp_active.pct")]
    stall = sorted(stall, key=lambda x: -float(x[1].replace(",", "") or 0))[:8]
    for k, v in stall:
        print(f"   stall {k.split('issue_stalled_')[1].replace('_per_warp_active.pct',''):40s} {v}")
if nsrc:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    if rows:
        h = rows[0]
        try:
            ci = h.index("Source")
            si = [i for i, x in enumerate(h) if x.startswith("# Samples") or x == "Warp Stall Sampling (All Samples)"]
            si = si[0] if si else None
        except ValueError:
            si = None
        if si is not None:
            body = [r for r in rows[1:] if len(r) > si and r[si].replace(",", "").isdigit()]
            tot = sum(int(r[si].replace(",", "")) for r in body) or 1
            body.sort(key=lambda r: -int(r[si].replace(",", "")))
            print("--- top source lines by stall samples")
            for r in body[:nsrc]:
                print(f"{100*int(r[si].replace(',',''))/tot:5.1f}%  {r[ci][:150]}")
