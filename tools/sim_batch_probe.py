"""What would launching k_similarity for several frames at once buy?  Emulated with k copies of the 360 M1 templates
in one class: the grid grows k-fold on the same linear memories."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
from _fixtures import load_case, load_templates
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
for k in (1, 2, 4, 8):
    det = Detector(nf, (4, 8))
    det.setClassTemplates(cid, np.concatenate([tdesc] * k), np.concatenate([feats] * k))
    det.match_records(case["img"], 90.0, [cid])
    ms = min(det.time_similarity(90.0, 20) for _ in range(3))
    print("copies %d: %.2f us per launch, %.2f us per 360 templates" % (k, ms * 1e3, ms * 1e3 / k))
    det.close()
