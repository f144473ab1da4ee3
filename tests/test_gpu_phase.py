"""Exhaustive: cv::phase (degrees) and the 16-bin orientation label for EVERY Sobel response the front end can see --
(dx, dy) in [-1020, 1020]^2, 4.2 M pairs -- as the orientation kernel's own device functions compute them
(sbm_phase_degrees) against the OpenCV restatement pinned to cv2 4.13 (oracle/shim fastAtan32f FMA form, convertTo's
round-half-even): angles bit for bit, labels identical.  The training path stores the angle itself as Feature::theta
(line2Dup.cpp:498-499), so a last-ulp difference there would surface in addTemplate_rotate's labels."""
import ctypes as C

import numpy as np
import pytest

from oracle import ref

pytestmark = pytest.mark.gpu


def all_pairs():
    v = np.arange(-1020, 1021, dtype=np.int32)
    dx, dy = np.meshgrid(v, v)
    return np.ascontiguousarray(dy.ravel()), np.ascontiguousarray(dx.ravel())


def reference_phase(dy, dx):
    L = ref.lib()
    x, y = dx.astype(np.float32), dy.astype(np.float32)
    ang = np.zeros(x.size, np.float32)
    assert L.ref_cv_phase(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), x.size, ang.ctypes.data_as(C.c_void_p)) == 0
    q = np.zeros(x.size, np.uint8)
    assert L.ref_cv_convert_u8(ang.ctypes.data_as(C.c_void_p), x.size, C.c_double(16.0 / 360.0), q.ctypes.data_as(C.c_void_p)) == 0
    return ang, q & 7


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
def test_phase_and_label_of_every_sobel_response():
    from shape_based_matching_b200 import _capi

    dy, dx = all_pairs()
    ang = np.zeros(dy.size, np.float32)
    lab = np.zeros(dy.size, np.uint8)
    _capi.check(_capi.lib().sbm_phase_degrees(0, dy.ctypes.data_as(C.c_void_p), dx.ctypes.data_as(C.c_void_p), dy.size,
                                              ang.ctypes.data_as(C.c_void_p), lab.ctypes.data_as(C.c_void_p)))
    want_ang, want_lab = reference_phase(dy, dx)
    bad = np.flatnonzero(lab != want_lab)
    assert bad.size == 0, "labels differ at %d pairs, first (dy, dx) = %s" % (bad.size, [(int(dy[i]), int(dx[i])) for i in bad[:5]])
    bad = np.flatnonzero(ang.view(np.uint32) != want_ang.view(np.uint32))
    assert bad.size == 0, "angles differ at %d pairs, first (dy, dx, got, want) = %s" % (
        bad.size, [(int(dy[i]), int(dx[i]), float(ang[i]), float(want_ang[i])) for i in bad[:5]])
