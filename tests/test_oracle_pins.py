"""CPU tests of the ORACLE: (1) the OpenCV restatements in oracle/shim against REAL cv2 4.13 outputs
(tests/golden/cv_pins.npz), (2) the reference's own code (oracle/_ref) against the committed goldens,
(3) the reference's YAML reader/writer through include/sbm/cv_compat.hpp.

Skipped when oracle/_ref/libsbm_ref.so is absent (it is built where /root/reference exists and travels
with the snapshot; it is git-ignored)."""
import ctypes as C
import os

import numpy as np
import pytest

from _fixtures import GOLDEN, load_case, load_templates
from oracle import ref

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(scope="module")
def pins():
    return dict(np.load(os.path.join(GOLDEN, "cv_pins.npz")))


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a):
    a = np.ascontiguousarray(a)
    if a.ndim == 2:
        a = a[:, :, None]
    return a, a.shape[1], a.shape[0], a.shape[2]


@pytest.mark.parametrize("name", ["g", "c", "s"])
def test_gaussian_sobel_pyrdown_match_cv2(pins, name):
    L = ref.lib()
    im, w, h, c = _img(pins[name])
    out = np.zeros_like(im)
    assert L.ref_cv_gaussian7(_vp(im), w, h, c, _vp(out)) == 0
    assert np.array_equal(out.reshape(pins[name + "_gauss"].shape), pins[name + "_gauss"])
    for dx, dy, key in ((1, 0, "_sobx"), (0, 1, "_soby")):
        o16 = np.zeros(im.shape, np.int16)
        assert L.ref_cv_sobel(_vp(im), w, h, c, dx, dy, _vp(o16)) == 0
        assert np.array_equal(o16.reshape(pins[name + key].shape), pins[name + key])
    want = pins[name + "_pyr"]
    od = np.zeros((h // 2, w // 2, c), np.uint8)
    assert L.ref_cv_pyrdown(_vp(im), w, h, c, _vp(od)) == 0
    assert np.array_equal(od.reshape(want.shape), want)


def test_pyrdown_odd_size(pins):
    L = ref.lib()
    im, w, h, c = _img(pins["odd"])
    od = np.zeros((h // 2, w // 2), np.uint8)
    assert L.ref_cv_pyrdown(_vp(im), w, h, c, _vp(od)) == 0
    assert np.array_equal(od, pins["odd_pyr"])


def test_phase_bit_exact(pins):
    L = ref.lib()
    x = np.ascontiguousarray(pins["phase_x"], np.float32)
    y = np.ascontiguousarray(pins["phase_y"], np.float32)
    out = np.zeros_like(x)
    assert L.ref_cv_phase(_vp(x), _vp(y), x.size, _vp(out)) == 0
    bad = np.flatnonzero(out.view(np.uint32) != pins["phase_deg"].view(np.uint32))
    assert bad.size == 0, "%d of %d angles differ, e.g. (%g,%g): %r vs %r" % (
        bad.size, x.size, x[bad[0]], y[bad[0]], out[bad[0]], pins["phase_deg"][bad[0]])


def test_convert_to_u8_rounds_half_even(pins):
    L = ref.lib()
    a = np.ascontiguousarray(pins["conv_in"], np.float32)
    out = np.zeros(a.size, np.uint8)
    assert L.ref_cv_convert_u8(_vp(a), a.size, C.c_double(16.0 / 360.0), _vp(out)) == 0
    assert np.array_equal(out, pins["conv_out"])


def test_erode_and_resize(pins):
    L = ref.lib()
    m = np.ascontiguousarray(pins["mask"])
    out = np.zeros_like(m)
    assert L.ref_cv_erode3(_vp(m), m.shape[1], m.shape[0], _vp(out)) == 0
    assert np.array_equal(out, pins["mask_erode"])
    for src_key, pre in (("c", "rs"), ("g", "rsg")):
        im, w, h, c = _img(pins[src_key])
        for i in range(4):
            f = float(pins["rs%d_f" % i])
            want = pins["%s%d" % (pre, i)]
            ow, oh = C.c_int(), C.c_int()
            buf = np.zeros(want.size + 64, np.uint8)
            assert L.ref_cv_resize_linear(_vp(im), w, h, c, C.c_double(f), C.c_double(f), _vp(buf), C.byref(ow), C.byref(oh)) == 0
            assert (oh.value, ow.value) == want.shape[:2]
            got = buf[: want.size].reshape(want.shape)
            # (training helper path: shapeInfo_producer::transform, line2Dup.h:379-405)
            assert np.array_equal(got, want)


@pytest.mark.parametrize("name", ["case1", "case2", "case0_1", "case0_2", "case0_3", "m1"])
def test_reference_reproduces_goldens(name):
    """oracle/_ref (the reference's own code) against the committed goldens -- whose quantised images were
    cross-checked against a REAL cv2 pipeline when they were generated (make_fixtures.py)."""
    case = load_case(name)
    cid, tdesc, feats, nf = load_templates(case["templates"])
    det = ref.RefDetector(nf, [int(t) for t in case["T"]])
    det.add_class(cid, tdesc, feats)
    thr = float(case["threshold"])
    got = det.match(case["img"], thr, [cid], omp_threads=1)
    want = case["matches"]
    assert len(got) == len(want)
    for k in ("x", "y", "template_id"):
        assert np.array_equal(got[k], want[k])
    assert np.array_equal(got["similarity"].view(np.uint32), want["similarity"].view(np.uint32))
    if name in ("case0_3", "case1"):
        for l in range(2):
            assert np.array_equal(det.front(case["img"], l)["quantized"], case["quantized%d" % l])


def test_multithreaded_reference_same_candidate_multiset():
    """the pre-unique candidate multiset is deterministic whatever the OpenMP thread count; only
    the final list varies.  (The parity definition of the GPU path relies on it.)"""
    case = load_case("case0_3")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    det = ref.RefDetector(nf, [4, 8])
    det.add_class(cid, tdesc, feats)
    want = np.sort(case["candidates"], order=["template_id", "x", "y"])
    got = np.sort(det.candidates(case["img"], 90.0, [cid]), order=["template_id", "x", "y"])
    assert np.array_equal(got, want)


def test_yaml_roundtrip_through_reference_reader(tmp_path):
    """writeClasses -> readClasses of the reference over cv_compat's FileStorage, plain and gzip."""
    cid, tdesc, feats, nf = load_templates("case0")
    a = ref.RefDetector(nf, [4, 8])
    a.add_class(cid, tdesc[:5], feats[: int(tdesc[:5, :, 4].sum())])
    for ext in ("yaml", "yaml.gz"):
        fmt = str(tmp_path / ("%s_templ." + ext))
        a.write_classes(fmt)
        b = ref.RefDetector(nf, [4, 8])
        assert b.read_class(cid, fmt) == 5
        td2, ft2 = b.export_class(cid)
        assert np.array_equal(td2, tdesc[:5]) and np.array_equal(ft2, feats[: ft2.shape[0]])
