"""Per-kernel table of the counters that matter here from an ncu report (`--set full`), read on the CPU box:

    python tools/ncu_summary.py gpurun_out/r2c_prof.ncu-rep [--json out.json]
"""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "us", 1e-3),
    ("smsp__inst_executed.sum", "Minst", 1e-6),
    ("sm__inst_executed_pipe_lsu.sum", "Mlsu", 1e-6),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
    ("dram__bytes_read.sum", "dramR_MB", 1e-6),
    ("dram__bytes_write.sum", "dramW_MB", 1e-6),
    ("lts__t_bytes.sum", "l2_MB", 1e-6),
    ("l1tex__t_bytes.sum", "l1_MB", 1e-6),
    ("l1tex__data_pipe_lsu_wavefronts.sum", "Mwavefronts", 1e-6),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%", 1),
    ("lts__t_sector_hit_rate.pct", "l2hit%", 1),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1pipe%", 1),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%", 1),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 1),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long", 1),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short", 1),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg", 1),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio", 1),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_bar", 1),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait", 1),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math", 1),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel", 1),
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    ix = {n: i for i, n in enumerate(hdr)}
    table = []
    for r in rows[2:]:
        d = {"kernel": r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("sbm::", "")}
        for k, name, sc in KEYS:
            if k in ix and r[ix[k]] not in ("", "n/a"):
                try:
                    d[name] = float(r[ix[k]].replace(",", "")) * sc
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
