"""Key metrics of every kernel in an .ncu-rep (ncu --set full), as a short text table for profiles/."""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
want = [("gpu__time_duration.sum", "time (us, under ncu)"), ("dram__bytes_read.sum", "dram read"),
        ("dram__bytes_write.sum", "dram write"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm % of peak"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor-pipe instructions"),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("launch__registers_per_thread", "registers / thread"), ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
        ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts")]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    units = dict(zip(hdr, rows[1]))
    print(f"== {d.get('Kernel Name', '')[:100]}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
    for k, label in want:
        if k in d and d[k] != "":
            print(f"   {label:28s} {d[k]} {units.get(k, '')}")
    st = {h.split("issue_stalled_")[1].split("_per")[0]: float(d[h]) for h in hdr
          if h.startswith("smsp__average_warps_issue_stalled") and "per_issue_active" in h and d[h]}
    top = sorted(st.items(), key=lambda x: -x[1])[:4]
    print("   top stalls (warps per issue): " + ", ".join(f"{k} {v:.2f}" for k, v in top))
