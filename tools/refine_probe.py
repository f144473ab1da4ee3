"""k_refine time against the number of coarse candidates (threshold sweep on the M1 frame): fixed chain or per-candidate work?"""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from _fixtures import load_case, load_templates
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
import shape_based_matching_b200._capi as capi
if len(sys.argv) > 1:
    capi.LIB_PATH = sys.argv[1]
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
for thr in (100.5, 99.9, 98.0, 95.0, 90.0, 85.0, 80.0, 70.0):
    det.match_records(case["img"], thr, [cid], cap=1 << 16)
    st = det.last_stats()
    t = det.time_kernels(thr, 20)
    print("thr %.1f coarse candidates %d: refine %.2f us, similarity %.2f us" % (thr, st["coarse_candidates"], t["refine_L0"] * 1e3, t["similarity"] * 1e3))
