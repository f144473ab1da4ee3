"""similarity_ms (CUDA events around k_similarity) as a function of the number of templates: separates the fixed
cost (launch + prologue latency) from the marginal cost per template."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
from _fixtures import load_case, load_templates
from shape_based_matching_b200 import Detector
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
d = torch.from_numpy(case["img"]).cuda()
for reps in (0, 1, 2, 4, 8):
    det = Detector(nf, (4, 8))
    if reps == 0:
        n1 = int(tdesc[:1, :, 4].sum())
        det.setClassTemplates("c0", tdesc[:1], feats[:n1])
    for r in range(reps):
        det.setClassTemplates("c%d" % r, tdesc, feats)
    det.set_timing(True)
    ts = []
    for i in range(40):
        det.match_device(d.data_ptr(), 1024, 1024, 3, 90.0)
        torch.cuda.synchronize()
        ts.append(det.last_timing())
    med = {k: float(np.median([t[k] for t in ts[5:]])) for k in ts[0]}
    print("templates=%4d" % max(1, 360 * reps), {k: round(v * 1e3, 1) for k, v in med.items()}, "us")
    det.close()
