"""Frames in flight against the frame size: is the device-resident rate at M1 bound by the kernels or by the rate at which
the GPU's front end starts graph nodes?  Same templates, same graph (10 nodes per frame), square frames of SIZE pixels
cut from / tiled out of the M1 frame.
    python tools/inflight_floor.py [lib]"""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
import shape_based_matching_b200._capi as capi
if len(sys.argv) > 1:
    capi.LIB_PATH = sys.argv[1]
from shape_based_matching_b200 import Detector
import bench
base, cid, tdesc, feats, nf = bench.load_workload()
B = 64
big = bench.make_frames(base, B, 1000)
for size in (256, 512, 768, 1024):
    frames = np.ascontiguousarray(big[:, :size, :size, :])
    det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
    d = torch.from_numpy(frames).cuda()
    recs = torch.zeros((B, 4096 * 20), dtype=torch.uint8, device="cuda"); cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream()
    def step():
        det.match_batch_device(d.data_ptr(), size * size * 3, B, size, size, 3, 90.0, [cid], recs.data_ptr(), 4096, cnt.data_ptr(), st.cuda_stream)
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(5): step()
    e1.record(st)
    torch.cuda.synchronize()
    print("%4d x %4d: %.1f us per frame in flight" % (size, size, e0.elapsed_time(e1) * 1e3 / (5 * B)))
    det.close()
