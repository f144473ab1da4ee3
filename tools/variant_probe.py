"""M1 on one GPU with a given build of the library: parity against the golden list, warm kernel times, frames in flight.

    python tools/variant_probe.py [lib.so] [--env K=V ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
args = [a for a in sys.argv[1:]]
for a in list(args):
    if "=" in a and not a.endswith(".so"):
        k, v = a.split("=", 1)
        os.environ[k] = v
        args.remove(a)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import shape_based_matching_b200._capi as capi  # noqa: E402
from _fixtures import load_case, load_templates  # noqa: E402

if args:
    capi.LIB_PATH = os.path.abspath(args[0])
from shape_based_matching_b200 import DEVICE_REC, Detector  # noqa: E402

case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
det = Detector(nf, (4, 8))
det.setClassTemplates(cid, tdesc, feats)
r = det.match_records(case["img"], 90.0, [cid])
g = case["matches"]
ok = len(r) == len(g) and np.array_equal(r["x"], g["x"]) and np.array_equal(r["similarity"].view(np.uint32), g["similarity"].view(np.uint32))
sim = [round(det.time_similarity(90.0, 50) * 1e3, 2) for _ in range(3)]
for _ in range(2):
    t = det.time_kernels(90.0, 30)
B = 64
frames = np.stack([np.roll(case["img"], (16 * i, 32 * i), axis=(0, 1)) for i in range(B)])
d = torch.from_numpy(frames).cuda()
recs = torch.zeros((B, 512 * DEVICE_REC.itemsize), dtype=torch.uint8, device="cuda")
cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
st = torch.cuda.Stream()
torch.cuda.set_stream(st)


def step():
    det.match_batch_device(d.data_ptr(), d[0].numel(), B, 1024, 1024, 3, 90.0, [cid], recs.data_ptr(), 512, cnt.data_ptr(), st.cuda_stream)


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(10):
    step()
e1.record(st)
torch.cuda.synchronize()
print(os.path.basename(capi.LIB_PATH), {k: v for k, v in os.environ.items() if k.startswith("SBM_")}, "golden" if ok else "MISMATCH",
      "layout", det.last_stats()["coarse_layout"], "| similarity us", sim, "| us/frame in flight %.2f" % (e0.elapsed_time(e1) / 10 / B * 1e3))
print("   ", {k: round(v * 1e3, 2) for k, v in t.items()}, "sum %.1f" % (1e3 * sum(v for k, v in t.items() if k != "frame_graph")))
