"""Condenses an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the handful of counters DESIGN.md and
bench.py's roofline block quote.  Usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/r01_x.txt"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print("# source:", path)
    for r in rows[2:]:
        print("kernel:", r[hdr.index("Kernel Name")])
        vals = {}
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                vals[w] = (r[i], units[i])
                print("  %-80s %s %s" % (w, r[i], units[i]))
        try:
            rd, wr = float(vals["dram__bytes_read.sum"][0]), float(vals["dram__bytes_write.sum"][0])
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
            tb = rd * scale[vals["dram__bytes_read.sum"][1]] + wr * scale[vals["dram__bytes_write.sum"][1]]
            t, tu = vals["gpu__time_duration.sum"]
            ts = float(t) * {"ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}[tu]
            print("  derived: dram traffic %.4f GB per launch, %.1f GB/s under ncu" % (tb / 1e9, tb / ts / 1e9))
            ls, lr = vals.get("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"), vals.get("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum")
            if ls and lr and float(lr[0]) > 0:
                print("  derived: global load sectors/request %.2f" % (float(ls[0]) / float(lr[0])))
        except Exception as e:
            print("  (derived values unavailable: %r)" % e)


if __name__ == "__main__":
    main(sys.argv[1])
