import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
from _fixtures import load_case, load_templates
import shape_based_matching_b200._capi as capi
if len(sys.argv) > 1:
    capi.LIB_PATH = sys.argv[1]
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
r = det.match_records(case["img"], 90.0, [cid])
for _ in range(2):
    t = det.time_kernels(90.0, 30)
print({k: round(v * 1e3, 2) for k, v in t.items()}, "us; sum of kernels %.1f" % (1e3 * sum(v for k, v in t.items() if k != "frame_graph")))
