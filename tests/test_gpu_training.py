"""Template training through the drop-in C++ classes (include/line2Dup.h via csrc/line2Dup_c.cpp; front end on the GPU)
against the reference's own Detector (oracle/_ref): BASELINE.md's M4 and M5 template sets are produced the way the
reference's callers produce them (shapeInfo_producer -> addTemplate -> addTemplate_rotate, test.cpp:262-330) and must come
out identical, feature by feature; and the produced M4 set finds the objects planted in the M4 frame like the reference."""
import numpy as np
import pytest

from oracle import ref
from shape_based_matching_b200 import line2dup
from shape_based_matching_b200 import workloads as wl

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")]


def test_m4_template_set_equals_reference_training():
    """8191-feature request (lowered in 5 % steps until Detector::match's 8192 assert passes, the same rule on both
    sides), 50 angles at 0.1 degree x scales {0.9, 1.0, 1.1}"""
    want_td, want_ft, want_infos = wl.m4_templates(trainer=ref.RefDetector(8191, (4, 8)), n_angles=50,
                                                   probe=lambda: ref.RefDetector(8191, (4, 8)))
    td, ft, infos = wl.m4_templates(n_angles=50)
    assert td.shape == (150, 2, 5)
    assert np.array_equal(infos.view(np.uint32), want_infos.view(np.uint32))
    assert np.array_equal(td, want_td), np.argwhere(td != want_td)[:5]
    assert np.array_equal(ft, want_ft)
    assert int(td[:, 0, 4].max()) > 7000 and int(td[:, :, 4].max()) < 8192


def test_m5_template_set_equals_reference_training():
    want_td, want_ft, _ = wl.m5_templates(trainer=ref.RefDetector(128, (4, 8)), n_angles=90)
    td, ft, _ = wl.m5_templates(n_angles=90)
    assert td.shape == (90, 2, 5) and np.array_equal(td, want_td) and np.array_equal(ft, want_ft)


def test_cxx_detector_theta_and_match_on_m4_frame():
    """features keep their theta through addTemplate_rotate (line2Dup.cpp:1395-1409), and the trained class, matched by the
    same C++ object, gives the reference's list on the M4 frame (templates around the planted 12.3 degree instance)"""
    from shape_based_matching_b200 import Detector

    obj, mask = wl.textured_object(wl.M4["seed"] + 1, wl.M4["object_size"], wl.M4["object_shapes"], wl.M4["object_max_r"])
    cxx = line2dup.CxxDetector(7000, (4, 8))
    rd = ref.RefDetector(7000, (4, 8))
    assert cxx.addTemplate(obj, "m4", mask, 7000) == 0 and rd.add_template(obj, "m4", mask, 7000) == 0
    centre = (obj.shape[1] / 2.0, obj.shape[0] / 2.0)
    for k in range(1, 9):
        th = np.float32(12.0) + np.float32(0.1) * k
        assert cxx.addTemplate_rotate("m4", 0, float(th), centre) == k
        assert rd.add_template_rotate("m4", 0, float(th), *centre) == k
    td, ft, th = cxx.export_class("m4", with_theta=True)
    wtd, wft, wth = rd.export_class("m4", with_theta=True)
    assert np.array_equal(td, wtd) and np.array_equal(ft, wft) and np.array_equal(th.view(np.uint32), wth.view(np.uint32))
    frame = wl.m4_frame()
    want = rd.match(frame, 90.0, ["m4"], omp_threads=1)
    got = cxx.match(frame, 90.0, ["m4"])
    assert len(want) >= 3 and len(got) == len(want)
    for k in ("x", "y", "template_id", "class_idx"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(got["similarity"].view(np.uint32), want["similarity"].view(np.uint32))
    # and through the C ABI mirror with the exported arrays
    det = Detector(7000, (4, 8))
    det.setClassTemplates("m4", td, ft)
    again = det.match_records(frame, 90.0, ["m4"])
    assert np.array_equal(again["x"], want["x"]) and np.array_equal(again["similarity"].view(np.uint32), want["similarity"].view(np.uint32))
