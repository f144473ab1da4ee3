"""Where the end-to-end M1 time goes: Detector.match_batch (Python wrapper + C call) against the C call alone and against
the H2D floor (the same pinned frames copied back to back on one stream).   python tools/e2e_breakdown.py"""
import ctypes as C, sys, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
import bench
from shape_based_matching_b200 import Detector
from shape_based_matching_b200._capi import MATCH_REC
base, cid, tdesc, feats, nf = bench.load_workload()
B = 256
frames = bench.make_frames(base, B, 1000)
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
pinned = torch.from_numpy(frames).pin_memory(); pin_np = pinned.numpy()
fl = [pin_np[i] for i in range(B)]
blk = pin_np
det.match_batch(fl, 90.0, [cid])
def t(f, n=5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n / B * 1e6
print("Detector.match_batch          %.1f us per frame" % t(lambda: det.match_batch(fl, 90.0, [cid])))
ptrs = (C.c_void_p * B)(*[f.ctypes.data for f in fl])
arr, ncls, _k = det._class_args([cid])
cap = 256 * B; begin = (C.c_int * (B + 1))(); out = np.zeros(cap, MATCH_REC)
def raw():
    rc = det._L.sbm_match_batch(det._h, ptrs, None, B, 1024, 1024, 3, 0, C.c_float(90.0), arr, ncls, out.ctypes.data_as(C.c_void_p), cap, begin)
    assert rc == 0
print("Detector.match_batch, one block %.1f us per frame" % t(lambda: det.match_batch(blk, 90.0, [cid])))
print("sbm_match_batch (C call only)  %.1f us per frame" % t(raw))
d_sink = torch.empty((B, 1024, 1024, 3), dtype=torch.uint8, device="cuda")
def floor():
    for i in range(B): d_sink[i].copy_(pinned[i], non_blocking=True)
print("H2D floor, one stream          %.1f us per frame" % min(t(floor, 1) for _ in range(3)))
streams = [torch.cuda.Stream() for _ in range(8)]
def floor8():
    for i in range(B):
        with torch.cuda.stream(streams[i % 8]): d_sink[i].copy_(pinned[i], non_blocking=True)
print("H2D floor, 8 streams in turn   %.1f us per frame" % min(t(floor8, 1) for _ in range(3)))
