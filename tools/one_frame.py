"""Two blocking matches of the M1 frame (the second is the one to profile: `ncu --launch-skip 7 --launch-count 7`)."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from _fixtures import load_case, load_templates
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
for _ in range(2):
    r = det.match_records(case["img"], 90.0, [cid])
print(len(r), det.last_stats())
