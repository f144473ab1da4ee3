"""profiles/<tag>_summary.md out of the round's captures (bench line, ncu tables, stage table, in-flight shares, timeline):
    python tools/profiles_summary.py r4 r4k"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, cap = sys.argv[1], sys.argv[2]
P = lambda n: os.path.join(ROOT, "profiles", n)
d = json.loads(open(P(cap + "_bench.json")).read())
kw = d["roofline"]["kernel_ms_warm"]
ncu = json.load(open(P(cap + "_ncu_frame.json")))
rows = "\n".join("| `%s` | %.2f | %.2f | %.1f | %.1f | %d | %.2f / %.2f | %.2f | %.2f | %.2f | %.2f |" % (
    k["kernel"], k["us"], k["Minst"], k["issue%"], k["warps%"], k["regs"], k["dramR_MB"], k["dramW_MB"], k["st_long"], k["st_bar"], k["st_math"], k["st_notsel"]) for k in ncu)
shares, lines, tl = (open(P(cap + n)).read() for n in ("_inflight_shares.txt", "_k_orient_mma_lines.txt", "_timeline.txt"))
head = open(P(tag + "_summary.md")).read().split("## Bench line (N = 1)")[0] if os.path.exists(P(tag + "_summary.md")) else "# %s\n\n" % tag
txt = head + f"""## Bench line (N = 1)

value {d['value']:.4e} evaluations/s = {d['ms_per_frame']*1e3:.2f} us per M1 frame (256 frames per step, 8 in flight; round 2 began at 20.7 us, round 1 ended
at 23.4); e2e {d['e2e']['ms_per_frame']*1e3:.1f} us per frame from pinned host memory to final host lists ({d['e2e']['pcie_floor']['e2e_over_floor']:.3f} x the H2D floor of this box,
{d['e2e']['pcie_floor']['h2d_us_per_frame']:.1f} us); one blocking `match` call {d['latency']['match_call_ms_host_to_host']*1e3:.0f} us; roofline (K5a, L2-resident working set): {d['roofline']['achieved']/1e3:.2f} TB/s of
{d['roofline']['peak']/1e3:.2f} TB/s measured L2 read peak = {d['roofline']['frac']:.3f}, DRAM traffic per launch {d['roofline']['traffic']} B against 71.86 MB algorithmic;
configuration 4: {d['c4_template_sharded']['ms_per_frame']:.2f} ms per frame; configuration 5: {d['c5_frame_sharded']['frames_per_s']:.0f} frames/s; the reference's CPU code
{d['cpu_baseline']['ms_per_frame']:.1f} ms per frame on {d['cpu_baseline']['cores']} host threads; clocks {d['clocks']['sm_mhz']:.0f} / {d['clocks']['sm_max_mhz']:.0f} MHz, no throttle reason.

Kernels alone, warm (us): """ + ", ".join("%s %.2f" % (k, v * 1e3) for k, v in kw.items()) + f"""

## ncu `--set full`, one M1 frame (cold-cache, serialised launches)

| kernel | us | M warp instr | issue-active % | warps active % | regs | DRAM MB read / written | stall long sb | barrier | math pipe | not selected |
|---|---|---|---|---|---|---|---|---|---|---|
{rows}

## `k_orient_mma<3,5>` by stage

```
{lines}```

## What each kernel costs with frames in flight

```
{shares}```

## Timeline of 64 frames in flight

```
{tl}```
"""
open(P(tag + "_summary.md"), "w").write(txt)
print("wrote", P(tag + "_summary.md"))
