"""How long does the HOST need to enqueue one device-resident frame (sbm_match_batch_device)?  If this is close to
the measured us/frame of bench.py's `value`, the device-resident throughput is launch-bound, not GPU-bound."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from bench import load_workload, make_frames, THRESHOLD, W, H, C, REC_CAP
from shape_based_matching_b200 import DEVICE_REC, Detector

base, cid, tdesc, feats, nf = load_workload()
B = 64
frames = make_frames(base, B, 1000)
det = Detector(nf, (4, 8), 30.0, 60.0)
det.setClassTemplates(cid, tdesc, feats)
d_frames = torch.from_numpy(frames).cuda()
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
d_recs = torch.zeros((B, REC_CAP * DEVICE_REC.itemsize), dtype=torch.uint8, device="cuda")
d_cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
def step():
    det.match_batch_device(d_frames.data_ptr(), W * H * C, B, W, H, C, THRESHOLD, [cid], d_recs.data_ptr(), REC_CAP, d_cnt.data_ptr(), stream.cuda_stream)
for _ in range(3):
    step()
torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter(); step(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print("enqueue %.1f us/frame, until done %.1f us/frame" % ((t1 - t0) / B * 1e6, (t2 - t0) / B * 1e6))
