"""Where a CTA of k_similarity / k_refine spends its time (experiment; needs a library built with -DSBM_TRACE):

    nvcc ... -DSBM_TRACE ... -o _variants/libsbm_trace.so ; python tools/stage_trace.py _variants/libsbm_trace.so

Thread 0 of every CTA adds its SM clock (cycles since CTA entry) at a few stages; printed: the mean over CTAs."""
import ctypes
import sys

sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from _fixtures import load_case, load_templates
import shape_based_matching_b200._capi as capi

capi.LIB_PATH = sys.argv[1] if len(sys.argv) > 1 else "_variants/libsbm_trace.so"
from shape_based_matching_b200 import Detector

case = load_case("m1")
cid, tdesc, feats, nf = load_templates("case1[:360]")
det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
L = capi.lib()
L.sbm_trace_read.argtypes = [ctypes.POINTER(ctypes.c_ulonglong), ctypes.c_int]
buf = (ctypes.c_ulonglong * 16)()
SIM = ("ctas", "staged (offsets + rawmin + barrier)", "gather loop done", "-", "-", "-", "-", "ctas skipped (tile beyond tp)")
REF = ("candidates", "count + candidate loaded", "barrier 1", "feature bases staged + barrier", "gather loop done (warp 0)",
       "partials barrier", "arg-max barrier", "record written")
for thr in (99.9, 90.0, 80.0):
    det.match_records(case["img"], thr, [cid], cap=1 << 16)
    st = det.last_stats()
    L.sbm_trace_read(buf, 1)
    t = det.time_kernels(thr, 10)
    L.sbm_trace_read(buf, 1)
    v = list(buf)
    print("threshold %.1f: %d coarse candidates; similarity %.2f us, refine %.2f us" % (thr, st["coarse_candidates"], t["similarity"] * 1e3, t["refine_L0"] * 1e3))
    for k, names in ((0, SIM), (1, REF)):
        n = max(v[8 * k], 1)
        print("  %s: %d CTA passes" % ("k_similarity" if k == 0 else "k_refine", v[8 * k]))
        for i in range(1, 8):
            if names[i] != "-":
                val = v[8 * k + i]
                print("    %-44s %s" % (names[i], ("%8.0f cycles" % (val / n)) if "skipped" not in names[i] else str(val)))
