"""Summarise .ncu-rep files (read with `ncu -i`, no GPU needed) into a small TSV for profiles/."""
import csv, subprocess, sys
WANT = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
out = open(sys.argv[1], "w")
first = True
for rep in sys.argv[2:]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(w) for w in WANT if w in hdr]
    if first:
        out.write("report\t" + "\t".join(f"{hdr[i]} [{units[i]}]" for i in idx) + "\n")
        first = False
    for r in rows[2:]:
        out.write(rep.split("/")[-1] + "\t" + "\t".join(r[i] for i in idx) + "\n")
