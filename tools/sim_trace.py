"""clock64() of one warp of k_similarity_bits at its phase boundaries (a -DK5B_TRACE build: tools/build_variant.py trace -DK5B_TRACE):
    python tools/sim_trace.py _variants/trace.so [n_templates]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
os.environ["SBM_BITS_PAIR"] = "0"
os.environ["SBM_BITS_SPLIT"] = "1"
import numpy as np  # noqa: E402

import shape_based_matching_b200._capi as capi  # noqa: E402
from _fixtures import load_case, load_templates  # noqa: E402

capi.LIB_PATH = os.path.abspath(sys.argv[1])
from shape_based_matching_b200 import Detector  # noqa: E402

n = int(sys.argv[2]) if len(sys.argv) > 2 else 360
case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:%d]" % n)
det = Detector(nf, (4, 8))
det.setClassTemplates(cid, tdesc, feats)
det.match_records(case["img"], 90.0, [cid])
us = det.time_similarity(90.0, 50) * 1e3
lib = ctypes.CDLL(capi.LIB_PATH)
buf = (ctypes.c_longlong * 16)()
assert lib.sbm_debug_k5b_trace(buf) == 0
t = list(buf)[:6]
names = ["entry", "item loaded", "offsets staged + barrier", "first group consumed", "main loop done", "finish done"]
print("templates %d  kernel %.2f us  features of the traced template's coarse level: %d" % (n, us, int(tdesc[n // 2, 1, 4])))
for i in range(1, 6):
    print("  %-28s +%6d cycles  (%.2f us at 1.965 GHz)" % (names[i], t[i] - t[i - 1], (t[i] - t[i - 1]) / 1965.0))
print("  total %d cycles = %.2f us" % (t[5] - t[0], (t[5] - t[0]) / 1965.0))
