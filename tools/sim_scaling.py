"""Coarse similarity kernel alone (50 launches in one graph) against the number of active templates, M1 frame:
is it one CTA's latency (flat) or throughput (linear)?   python tools/sim_scaling.py [lib.so] [K=V ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
args = sys.argv[1:]
for a in list(args):
    if "=" in a and not a.endswith(".so"):
        k, v = a.split("=", 1)
        os.environ[k] = v
        args.remove(a)
import numpy as np  # noqa: E402

import shape_based_matching_b200._capi as capi  # noqa: E402
from _fixtures import load_case, load_templates  # noqa: E402

if args:
    capi.LIB_PATH = os.path.abspath(args[0])
from shape_based_matching_b200 import Detector  # noqa: E402

case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
out = []
for n in (37, 74, 148, 222, 296, 360, 444, 592, 720, 1080, 1440):
    reps = (n + 359) // 360
    td = np.concatenate([tdesc] * reps)[:n]
    ft = np.concatenate([feats] * reps)[: int(td[:, :, 4].sum())]
    det = Detector(nf, (4, 8))
    det.setClassTemplates(cid, td, ft)
    det.match_records(case["img"], 90.0, [cid])
    us = min(det.time_similarity(90.0, 50) for _ in range(3)) * 1e3
    out.append("%d: %.2f" % (n, us))
    det.close()
print({k: v for k, v in os.environ.items() if k.startswith("SBM_")}, "templates: us |", "  ".join(out))
