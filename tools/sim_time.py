import sys, os, shutil
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
from _fixtures import load_case, load_templates
import shape_based_matching_b200._capi as capi
if len(sys.argv) > 1:
    capi.LIB_PATH = sys.argv[1]
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
r = det.match_records(case["img"], 90.0, [cid])
print(capi.LIB_PATH, "matches", len(r), "similarity us/launch:", [round(det.time_similarity(90.0, 50) * 1e3, 2) for _ in range(3)])
