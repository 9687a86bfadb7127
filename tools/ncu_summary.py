"""Summarise an .ncu-rep (read on the CPU box): one column per captured launch, key roofline metrics as rows."""
import csv, io, subprocess, sys
rep = sys.argv[1]
names = sys.argv[2].split(",") if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__cluster_dim_x", "sm__cycles_elapsed.max", "smsp__cycles_active.avg"]
print("metric," + ",".join(names if names else [r[hdr.index("Kernel Name")][:24] for r in data]))
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(f"{w} [{units[i]}]," + ",".join(r[i] for r in data))
