"""The plain C++ restatement (oracle/sbm_oracle.cpp) against the committed outputs of the reference's own code
(tests/golden/*.npz from oracle/_ref): every intermediate and both lists must be identical.  This pins the
restatement; the GPU parity tests then have two independent checkers."""
import hashlib

import numpy as np
import pytest

from _fixtures import load_case, load_templates, to_gray
from oracle import port


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def same_list(a, b):
    assert len(a) == len(b)
    for k in ("x", "y", "template_id"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))


@pytest.mark.parametrize("name", ["case0_3", "case0_2", "case2", "case1"])
def test_port_equals_reference_goldens(name):
    case = load_case(name)
    cid, tdesc, feats, nf = load_templates(case["templates"])
    det = port.PortDetector([int(t) for t in case["T"]], float(case["weak"]))
    det.add_class(tdesc, feats)
    det.process(case["img"])
    for l in range(2):
        q = det.get(l, 0).reshape(case["quantized%d" % l].shape)
        assert np.array_equal(q, case["quantized%d" % l])
        assert sha(det.get(l, 1)) == str(case["spread%d_sha" % l])
        assert sha(det.get(l, 2)) == str(case["response%d_sha" % l])
        assert sha(det.get(l, 3)) == str(case["linmem%d_sha" % l])
    for t, want in zip(case["sim_ids"], case["sim_maps"]):
        assert np.array_equal(det.similarity(0, int(t)), want)
    thr = float(case["threshold"])
    same_list(det.match(thr, finalize=False), case["candidates"])
    same_list(det.match(thr, finalize=True), case["matches"])


def test_port_gray_frame():
    case = load_case("case0_3")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    det = port.PortDetector([int(t) for t in case["T"]], float(case["weak"]))
    det.add_class(tdesc, feats)
    det.process(to_gray(case["img"]))
    for l in range(2):
        assert np.array_equal(det.get(l, 0).reshape(case["gray_quantized%d" % l].shape), case["gray_quantized%d" % l])
    same_list(det.match(float(case["threshold"])), case["gray_matches"])


def test_port_masked_and_two_classes():
    """the restatement against the reference's own outputs with a mask and with two classes in three request orders"""
    from _fixtures import load_masked_two_class

    z = load_masked_two_class()
    det = port.PortDetector([int(t) for t in z["T"]], float(z["weak"]))
    ids = sorted(c[0] for c in z["classes"])
    for cid in ids:  # class slots follow std::map order
        td, ft = next((c[1], c[2]) for c in z["classes"] if c[0] == cid)
        det.add_class(td, ft)
    thr = float(z["threshold"])
    for masked in (False, True):
        det.process(z["img"], z["mask"] if masked else None)
        pre = "masked_" if masked else ""
        for tag, req in z["requests"]:
            idx = tuple(ids.index(c) for c in req) if req else tuple(range(len(ids)))
            for kind, fin in (("candidates_", False), ("matches_", True)):
                got, want = det.match(thr, class_idx=idx, finalize=fin), z[pre + kind + tag]
                same_list(got, want)
                assert np.array_equal(got["class_idx"], want["class_idx"])
        if masked:
            for l in range(2):
                assert np.array_equal(det.get(l, 0).reshape(z["masked_quantized%d" % l].shape), z["masked_quantized%d" % l])
                assert sha(det.get(l, 3)) == str(z["masked_linmem%d_sha" % l])
            assert np.array_equal(det.similarity(ids.index("a_ring"), 7), z["masked_sim_ring7"])
