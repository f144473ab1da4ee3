"""Per-kernel table of the counters that matter here from an ncu report (`--set full`), read on the CPU box:

    python tools/ncu_summary.py gpurun_out/r2c_prof.ncu-rep [--json out.json]
"""
import csv
import io
import json
import subprocess
import sys

# (metric, short name, unit the table prints: time in us, data in MB, counts in millions where the name says so)
KEYS = [
    ("gpu__time_duration.sum", "us", "us"),
    ("smsp__inst_executed.sum", "Minst", "M"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", ""),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%", ""),
    ("launch__registers_per_thread", "regs", ""),
    ("launch__grid_size", "grid", ""),
    ("launch__block_size", "block", ""),
    ("dram__bytes_read.sum", "dramR_MB", "MB"),
    ("dram__bytes_write.sum", "dramW_MB", "MB"),
    ("lts__t_sectors.sum", "l2_Msectors", "M"),
    ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1_ld_Msectors", "M"),
    ("l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1_ld_Mwavefronts", "M"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%", ""),
    ("lts__t_sector_hit_rate.pct", "l2hit%", ""),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1pipe%", ""),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts%", ""),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", ""),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%", ""),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%", ""),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long", ""),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short", ""),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg", ""),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio", ""),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_bar", ""),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait", ""),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math", ""),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel", ""),
]
# ncu prints every column in a unit of its own choosing (second row of the raw page)
TO_BASE = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9,
           "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
WANT = {"us": 1e6, "MB": 1e-6, "M": 1e-6, "": 1.0}


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(hdr)}
    table = []
    for r in rows[2:]:
        d = {"kernel": r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("sbm::", "")}
        for k, name, want in KEYS:
            if k in ix and r[ix[k]] not in ("", "n/a"):
                try:
                    d[name] = float(r[ix[k]].replace(",", "")) * TO_BASE.get(units[ix[k]], 1.0) * WANT[want]
                except ValueError:
                    pass
        table.append(d)
    names = ["kernel"] + [n for _, n, _ in KEYS]
    for d in table:
        print(d["kernel"][:44].ljust(44), " ".join("%s=%s" % (n, ("%.2f" % d[n]) if n in d else "-") for n in names[1:]))
    if "--json" in sys.argv:
        json.dump(table, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
