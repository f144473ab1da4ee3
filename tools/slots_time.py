import sys, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
from _fixtures import load_case, load_templates
import shape_based_matching_b200._capi as capi
if len(sys.argv) > 1:
    capi.LIB_PATH = sys.argv[1]
from shape_based_matching_b200 import Detector, DEVICE_REC
sys.path.insert(0, ".")
import bench
base, cid, tdesc, feats, nf = bench.load_workload()
B = 64
frames = bench.make_frames(base, B, 1000)
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
d = torch.from_numpy(frames).cuda()
recs = torch.zeros((B, 4096 * 20), dtype=torch.uint8, device="cuda"); cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream()
def step():
    det.match_batch_device(d.data_ptr(), 1024 * 1024 * 3, B, 1024, 1024, 3, 90.0, [cid], recs.data_ptr(), 4096, cnt.data_ptr(), st.cuda_stream)
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record(st)
for _ in range(5): step()
e1.record(st); tc = time.perf_counter() - t0
torch.cuda.synchronize()
print("slots env", __import__("os").environ.get("SBM_DEVICE_SLOTS"), "device us/frame %.1f  cpu-enqueue us/frame %.1f" % (e0.elapsed_time(e1) * 1e3 / (5 * B), tc * 1e6 / (5 * B)))
def serial(stream):
    for i in range(B):
        det.match_device(d[i].data_ptr(), 1024, 1024, 3, 90.0, [cid], d_records=recs[i].data_ptr(), cap=4096, d_count=cnt[i].data_ptr(), stream=stream.cuda_stream)
for name, stream in (("default", torch.cuda.current_stream()), ("side", torch.cuda.Stream())):
    with torch.cuda.stream(stream):
        for _ in range(3): serial(stream)
        torch.cuda.synchronize()
        t0 = time.perf_counter(); e0.record(stream)
        for _ in range(5): serial(stream)
        e1.record(stream); tc = time.perf_counter() - t0
        torch.cuda.synchronize()
        print("serial match_device on", name, "stream: device us/frame %.1f cpu us/frame %.1f" % (e0.elapsed_time(e1) * 1e3 / (5 * B), tc * 1e6 / (5 * B)))
