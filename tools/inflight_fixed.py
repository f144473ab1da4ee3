"""How much of the M1 frame time in flight is a fixed cost per frame (kernel starts, tails, under-filled grids) and how much
is proportional to the pixels?  The M1 frame against two / four of them stacked into one 1024 x 2048 / 2048 x 2048 frame (same
templates: the similarity and refinement work also doubles / quadruples).
    python tools/inflight_fixed.py"""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
from shape_based_matching_b200 import Detector
import bench
base, cid, tdesc, feats, nf = bench.load_workload()
B = 32
big = bench.make_frames(base, 4 * B, 1000)
for ny, nx in ((1, 1), (2, 1), (2, 2)):
    k = ny * nx
    fr = big[: k * B].reshape(B, ny, nx, 1024, 1024, 3).transpose(0, 1, 3, 2, 4, 5).reshape(B, ny * 1024, nx * 1024, 3)
    frames = np.ascontiguousarray(fr)
    h, w = frames.shape[1:3]
    det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
    d = torch.from_numpy(frames).cuda()
    recs = torch.zeros((B, 16384 * 20), dtype=torch.uint8, device="cuda"); cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream()
    def step():
        det.match_batch_device(d.data_ptr(), h * w * 3, B, w, h, 3, 90.0, [cid], recs.data_ptr(), 16384, cnt.data_ptr(), st.cuda_stream)
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(5): step()
    e1.record(st)
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e3 / (5 * B)
    print("%d x %d (%d M1 frames per frame): %.1f us per frame in flight = %.1f us per M1 frame" % (w, h, k, t, t / k))
    det.close()
