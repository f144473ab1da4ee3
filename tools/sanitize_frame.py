"""One small colour frame and one gray frame through the whole pipeline (graph path and eager path with a mask), for
compute-sanitizer:  compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_frame.py"""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
from _fixtures import load_case, load_templates, to_gray
from shape_based_matching_b200 import Detector
case = load_case("case0_3")
cid, tdesc, feats, nf = load_templates(case["templates"])
for img in (case["img"], np.ascontiguousarray(to_gray(case["img"])[:, :, None])):
    det = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]), device=0)
    det.setClassTemplates(cid, tdesc, feats)
    r = det.match_records(img, float(case["threshold"]), [cid])
    mask = np.zeros(img.shape[:2], np.uint8); mask[32:200, 40:400] = 255
    rm = det.match_records(img, float(case["threshold"]), [cid], masks=mask)
    print(img.shape, len(r), len(rm))
    det.close()
