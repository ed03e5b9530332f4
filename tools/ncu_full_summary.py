"""Per-launch summary of an `ncu --set full` report:  python tools/ncu_full_summary.py X.ncu-rep"""
import csv, io, subprocess, sys
WANT = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
        "l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sector_hit_rate.pct",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__block_size", "launch__grid_size",
        "launch__registers_per_thread", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(h, r)); u = dict(zip(h, units))
    print(d["Kernel Name"])
    for k in WANT:
        if k in d: print("   %-85s %s %s" % (k, d[k], u[k]))
    f = lambda k: float(d[k].replace(",", ""))
    rd, wr, t = f("dram__bytes_read.sum"), f("dram__bytes_write.sum"), f("gpu__time_duration.sum")
    sect = f("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"); cyc = f("sm__cycles_elapsed.max")
    l2s = f("lts__t_sectors_srcunit_tex_op_read.sum")
    print("   -> DRAM traffic per launch %.1f MB (read %.1f + write %.1f); %.0f GB/s of DRAM traffic; "
          "%.2f global-load sectors/clk/SM; L2 -> SM %.0f MB = %.2f TB/s\n"
          % (rd + wr, rd, wr, (rd + wr) / t * 1e3, sect / cyc / 148, l2s * 32 / 1e6, l2s * 32 / 1e6 / t))
