"""Pulls the judged metrics out of an .ncu-rep (ncu -i rep --page raw --csv): per captured launch, duration, DRAM
bytes, tensor-pipe activity, issue/stall summary.  usage: ncu_raw_pick.py file.ncu-rep"""
import csv, io, subprocess, sys
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__inst_executed_pipe_tensor.sum",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max", "gpc__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "launch__grid_size",
        "smsp__cycles_active.avg", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "lts__t_sector_hit_rate.pct", "sm__clock_rate.avg", "gpc__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
ki = hdr.index("Kernel Name")
print("launches: %d" % len(data))
for r in data:
    print("kernel:", r[ki][:90], "grid", r[hdr.index("Grid Size")] if "Grid Size" in hdr else "")
for k in KEYS:
    if k in hdr:
        i = hdr.index(k)
        print("%-75s %-10s %s" % (k, units[i], [r[i] for r in data]))
