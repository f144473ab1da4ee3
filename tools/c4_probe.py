"""Configuration 4 in small: the M4 frame x the first N angles of every scale (default 100 -> 300 templates, 1/12 of M4),
coarse similarity kernel time and whole-frame device time for a given build of the library.

    python tools/c4_probe.py [lib.so] [K=V ...] [--angles N]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
args = list(sys.argv[1:])
n_angles = 100
if "--angles" in args:
    i = args.index("--angles")
    n_angles = int(args[i + 1])
    del args[i : i + 2]
for a in list(args):
    if "=" in a and not a.endswith(".so"):
        k, v = a.split("=", 1)
        os.environ[k] = v
        args.remove(a)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import shape_based_matching_b200._capi as capi  # noqa: E402

if args:
    capi.LIB_PATH = os.path.abspath(args[0])
from shape_based_matching_b200 import Detector  # noqa: E402
from shape_based_matching_b200 import workloads as wl  # noqa: E402

cache = os.path.join(ROOT, "gpurun_out", "_m4_%d.npz" % n_angles)
if os.path.exists(cache):
    z = np.load(cache)
    tdesc, feats, frame = z["tdesc"], z["feats"], z["frame"]
else:
    tdesc, feats, _ = wl.m4_templates(n_angles=n_angles)
    frame = wl.m4_frame()
    try:
        np.savez(cache, tdesc=tdesc, feats=feats, frame=frame)
    except Exception:
        pass
det = Detector(8191, (4, 8))
det.setClassTemplates("m4", tdesc, feats)
r = det.match_records(frame, 90.0, ["m4"], cap=1 << 16)
st = det.last_stats()
sim = [det.time_similarity(90.0, 3) for _ in range(2)]
t = det.time_kernels(90.0, 3)
sig = int(np.sum(r["x"].astype(np.int64) * 7 + r["y"] * 3 + r["template_id"])) if len(r) else 0
print(os.path.basename(capi.LIB_PATH), {k: v for k, v in os.environ.items() if k.startswith("SBM_")}, "templates", tdesc.shape[0], "matches", len(r), "sig", sig,
      "| similarity ms", [round(v, 3) for v in sim], "= %.1f T evals/s" % (st["coarse_evals"] / (min(sim) * 1e-3) / 1e12),
      "| frame_graph ms %.3f" % t["frame_graph"])
