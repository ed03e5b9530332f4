"""Per-launch table of an `ncu --set full` report: python tools/ncu_kernels.py X.ncu-rep"""
import csv, io, subprocess, sys
WANT = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("lts__t_sectors_srcunit_tex_op_read.sum", "L2rdsect"), ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "ldsect"),
        ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_bar"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
        ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio"),
        ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "st_membar"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
        ("smsp__inst_executed.sum", "inst"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum", "smemconfl"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "smemwf")]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(h, r)); u = dict(zip(h, units))
    print(d["Kernel Name"][:70])
    line = []
    for k, nm in WANT:
        if k in d:
            try:
                v = float(d[k].replace(",", ""))
                if u[k] == "byte": v /= 1e6
                if u[k] == "Kbyte": v /= 1e3
                if u[k] == "Gbyte": v *= 1e3
                if u[k] == "ns": v /= 1e3
                if u[k] == "ms": v *= 1e3
                line.append("%s=%.4g" % (nm, v))
            except ValueError:
                line.append("%s=%s" % (nm, d[k]))
    print("    " + "  ".join(line))
