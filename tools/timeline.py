"""A timeline of M1 frames in flight without a profiler: a -DSBM_TRACE build (tools/build_variant.py trace -DSBM_TRACE) lets
every CTA of the frame kernels record {global timer at start, at end, SM, kernel}; this reads the records of one batch of 64
frames and prints, per kernel: launches, CTAs, mean CTA duration, the span of one launch (first CTA start -> last CTA end) alone
in the batch; and for the GPU: how many SMs hold at least one CTA over time, resident CTAs, and which kernels run together.
    python tools/build_variant.py trace -DSBM_TRACE && python tools/timeline.py _variants/trace.so"""
import ctypes as C
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
import shape_based_matching_b200._capi as capi
capi.LIB_PATH = sys.argv[1]
from shape_based_matching_b200 import Detector
import bench
NAMES = {0: "orient", 1: "pyrdown", 2: "spread", 3: "similarity", 4: "refine", 5: "export"}
base, cid, tdesc, feats, nf = bench.load_workload()
B = 64
frames = bench.make_frames(base, B, 1000)
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
d = torch.from_numpy(frames).cuda()
recs = torch.zeros((B, 4096 * 20), dtype=torch.uint8, device="cuda"); cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream()
def step():
    det.match_batch_device(d.data_ptr(), 1024 * 1024 * 3, B, 1024, 1024, 3, 90.0, [cid], recs.data_ptr(), 4096, cnt.data_ptr(), st.cuda_stream)
lib = capi.lib()
lib.sbm_debug_timeline.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_uint)]
for _ in range(3): step()
torch.cuda.synchronize()
n = C.c_uint(0)
lib.sbm_debug_timeline(None, 0, C.byref(n))  # reset
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st); step(); e1.record(st)
torch.cuda.synchronize()
print("batch of %d frames: %.1f us per frame (events)" % (B, e0.elapsed_time(e1) * 1e3 / B))
dt = np.dtype([("t0", "<u8"), ("t1", "<u8"), ("kid", "<u4"), ("smid", "<u4"), ("cta", "<u4"), ("grid", "<u4")])
buf = np.zeros(1 << 20, dtype=dt)
lib.sbm_debug_timeline(buf.ctypes.data_as(C.c_void_p), len(buf), C.byref(n))
r = buf[: n.value]
T0 = r["t0"].min()
t0 = (r["t0"] - T0).astype(np.float64) * 1e-3
t1 = (r["t1"] - T0).astype(np.float64) * 1e-3
span = t1.max()
print("%d CTA records, span %.1f us = %.1f us per frame; timer steps seen: %s ns" % (len(r), span, span / B, np.unique(np.diff(np.sort(r["t0"])))[:4]))
key = r["kid"].astype(np.int64) * 100000 + r["grid"]
print("%-22s %8s %9s %12s %14s %16s" % ("kernel (grid)", "CTAs", "launches", "CTA us mean", "CTA us p90", "CTA-us per frame"))
for k in np.unique(key):
    m = key == k
    dur = t1[m] - t0[m]
    g = int(k % 100000)
    print("%-22s %8d %9.1f %12.2f %14.2f %16.1f" % ("%s (%d)" % (NAMES[int(k // 100000)], g), m.sum(), m.sum() / g, dur.mean(), np.percentile(dur, 90), dur.sum() / B))
m = (r["kid"] == 0) & (r["grid"] == r["grid"][r["kid"] == 0].max())
if m.any() and r["cta"][m].max() > 0:  # `cta` holds ns from a CTA's start to SBM_TRACE_MID (k_orient_mma: source window staged)
    mid = r["cta"][m].astype(np.float64) * 1e-3
    print("orient level 0: a CTA waits %.2f us (p90 %.2f) for its staged window, of %.2f us alive" % (mid.mean(), np.percentile(mid, 90), (t1[m] - t0[m]).mean()))
# occupancy over time
nsm = int(r["smid"].max()) + 1
step_us = 0.25
nb = int(span / step_us) + 2
cta_res = np.zeros(nb)
sm_busy = np.zeros((nsm, nb), dtype=bool)
kern_active = np.zeros((6, nb), dtype=bool)
for a, b_, sm, kid in zip(t0, t1, r["smid"], r["kid"]):
    i0, i1 = int(a / step_us), int(b_ / step_us) + 1
    cta_res[i0:i1] += 1
    sm_busy[sm, i0:i1] = True
    kern_active[kid, i0:i1] = True
busy = sm_busy.sum(0)
print("SMs holding at least one CTA: mean %.1f of %d; time with < 74 SMs busy: %.1f %%, with 0: %.1f %%" % (busy.mean(), nsm, 100 * (busy < nsm / 2).mean(), 100 * (busy == 0).mean()))
print("resident CTAs: mean %.0f; kernels (kinds) active at once: mean %.2f" % (cta_res.mean(), kern_active.sum(0).mean()))
hist = np.bincount(kern_active.sum(0), minlength=7)
print("time share with k kinds of kernels active, k = 0..6:", np.round(100 * hist / hist.sum(), 1))
for kid in range(6):
    print("  %-10s active %.1f %% of the time" % (NAMES[kid], 100 * kern_active[kid].mean()))
# registers / thread slots held per SM over time (registers per CTA from -Xptxas -v, rounded up as the hardware allocates them)
REGS = {0: 64 * 256, 1: 72 * 128, 2: 32 * 256, 3: 128 * 128, 4: 72 * 128, 5: 32 * 256}
THREADS = {0: 256, 1: 128, 2: 256, 3: 128, 4: 128, 5: 256}
R = np.zeros((nsm, nb)); TH = np.zeros((nsm, nb))
for a, b_, sm, kid in zip(t0, t1, r["smid"], r["kid"]):
    i0, i1 = int(a / step_us), int(b_ / step_us) + 1
    R[sm, i0:i1] += REGS[int(kid)]; TH[sm, i0:i1] += THREADS[int(kid)]
core = slice(int(100 / step_us), int((span - 100) / step_us))  # steady state: 100 us trimmed at both ends
print("steady state: registers held per SM mean %.0f of 65536 (%.0f %%), thread slots %.0f of 2048 (%.0f %%), CTAs %.2f" % (
    R[:, core].mean(), 100 * R[:, core].mean() / 65536, TH[:, core].mean(), 100 * TH[:, core].mean() / 2048, (cta_res[core] / nsm).mean()))
print("share of (SM, time) cells without room for another 16 K-register CTA: %.1f %%" % (100 * (R[:, core] > 65536 - 16384).mean()))
print("register x time per frame (M register-us):")
for k in np.unique(key):
    m = key == k
    print("  %-22s %6.1f" % ("%s (%d)" % (NAMES[int(k // 100000)], int(k % 100000)), REGS[int(k // 100000)] * (t1[m] - t0[m]).sum() / B / 1e6))
print("  capacity: %.1f" % (65536 * nsm * span / B / 1e6))
np.save("gpurun_out/timeline_records.npy", r)
