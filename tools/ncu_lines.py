"""Warp instructions and stall samples per source line of one kernel of an ncu report captured with --import-source on:
    python tools/ncu_lines.py report.ncu-rep 'k_orient<3, 32' [top_n] [lo:hi:name ...]
lo:hi:name groups source lines lo..hi of sbm_kernels.cuh under a name (the stage table in DESIGN.md)."""
import collections
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 and sys.argv[3].isdigit() else 20
groups = [a.split(":") for a in sys.argv[3:] if a.count(":") == 2]
base = kern.split("<")[0]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + base, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
# one section per (source file, function): "File Path", "Function Name", a header row ("Line No", ...), then source lines, each
# followed by the SASS attributed to it.  An instruction inlined from a helper is listed under the helper's line AND under
# the call site: it is counted once, at the LAST line of sbm_kernels.cuh that lists it (helpers are defined above the kernels
# that call them, so that is the call site inside the kernel).
at = {}
for i, r in enumerate(rows):
    if r and r[0] == "File Path" and r[1].endswith("sbm_kernels.cuh") and kern in rows[i + 1][1]:
        hdr = rows[i + 2]
        iex, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        line = None
        for r2 in rows[i + 3 :]:
            if r2 and r2[0] == "File Path":
                break
            if r2[0] != "":
                line = int(r2[0])
                continue
            try:
                at[r2[2]] = (line, int(r2[iex]), int(r2[ismp]), r2[3])
            except ValueError:
                pass
        break
by, smp, ops = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
for line, n, sm, text in at.values():
    by[line] += n
    smp[line] += sm
    w = text.split()
    ops[line][(w[1] if w[0].startswith("@") else w[0]).split(".")[0]] += n
tot, ts = sum(by.values()), max(1, sum(smp.values()))
print("kernel %s: %d warp instructions attributed, %d stall samples" % (kern, tot, ts))
if groups:
    g = collections.Counter()
    gs = collections.Counter()
    for l, n in by.items():
        name = next((nm for lo, hi, nm in groups if int(lo) <= l <= int(hi)), "other")
        g[name] += n
        gs[name] += smp[l]
    for k, v in sorted(g.items(), key=lambda x: -x[1]):
        print("  %-28s %9d  %5.1f %% of instructions  %5.1f %% of samples" % (k, v, 100.0 * v / tot, 100.0 * gs[k] / ts))
for l, n in by.most_common(top):
    print("  line %4d %9d %5.1f %%  %s" % (l, n, 100.0 * n / tot, dict(ops[l].most_common(5))))
