"""ctypes view of the drop-in C++ classes (include/line2Dup.h, libline2Dup_b200.so) through csrc/line2Dup_c.cpp:
the host code a C++ caller links -- training (addTemplate: GPU front end + host feature selection,
addTemplate_rotate), YAML persistence, match, setGroup, shapeInfo_producer -- driven from Python.

``detector.Detector`` talks to the C ABI directly and is what the benchmarks time; this module exists so that template
sets are PRODUCED the way the reference's callers produce them (/root/reference/test.cpp:262-330: shapeInfo_producer ->
addTemplate -> addTemplate_rotate) and so that the C++ layer is tested from the same Python tests as everything else.
"""
import ctypes as C
import os

import numpy as np

from ._capi import MATCH_REC, PKG

HOST_LIB_PATH = os.path.join(PKG, "libline2Dup_b200.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise ImportError("%s is missing: build it with `python -m shape_based_matching_b200.build`" % HOST_LIB_PATH)
        L = C.CDLL(HOST_LIB_PATH)
        L.l2d_last_error.restype = C.c_char_p
        L.l2d_create.restype = C.c_void_p
        L.l2d_create.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int, C.c_float, C.c_float, C.c_int]
        L.l2d_destroy.argtypes = [C.c_void_p]
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.l2d_add_template.argtypes = [vp, vp, ci, ci, ci, vp, C.c_char_p, ci]
        L.l2d_add_template_rotate.argtypes = [vp, C.c_char_p, ci, cf, cf, cf]
        L.l2d_num_templates.argtypes = [vp, C.c_char_p]
        L.l2d_class_sizes.argtypes = [vp, C.c_char_p, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
        L.l2d_export_class.argtypes = [vp, C.c_char_p, vp, vp, vp]
        L.l2d_read_class.argtypes = [vp, C.c_char_p, C.c_char_p]
        L.l2d_write_classes.argtypes = [vp, C.c_char_p]
        L.l2d_match.argtypes = [vp, vp, ci, ci, ci, vp, cf, C.c_char_p, vp, ci, C.POINTER(ci)]
        L.l2d_group_unique_id.argtypes = [vp]
        L.l2d_set_group.argtypes = [vp, ci, ci, vp, ci]
        L.l2d_shape_infos.argtypes = [vp, ci, vp, ci, cf, cf, vp, ci, C.POINTER(ci)]
        L.l2d_shape_transform.argtypes = [vp, ci, ci, ci, cf, cf, ci, vp, C.POINTER(ci), C.POINTER(ci)]
        _lib = L
    return _lib


def _img_args(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, c = img.shape
    return img, (img.ctypes.data_as(C.c_void_p), w, h, c)


class CxxDetector:
    """line2Dup::Detector of include/line2Dup.h (same calls as the reference's class, /root/reference/line2Dup.h:257-333)"""

    def __init__(self, num_features=63, T=(4, 8), weak=30.0, strong=60.0, device=0):
        self.L = lib()
        arr = (C.c_int * len(T))(*[int(t) for t in T])
        self.h = C.c_void_p(self.L.l2d_create(int(num_features), arr, len(T), weak, strong, int(device)))
        if not self.h:
            raise RuntimeError(self.L.l2d_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.l2d_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc < 0:
            raise RuntimeError(self.L.l2d_last_error().decode())
        return rc

    def addTemplate(self, img, class_id, mask=None, num_features=0):
        """-> template id, or -1 when the image yields too few features (line2Dup.cpp:514-517, 1342-1343)"""
        img, a = _img_args(img)
        m = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            assert mask.shape == img.shape[:2]
            m = mask.ctypes.data_as(C.c_void_p)
        rc = self.L.l2d_add_template(self.h, *a, m, class_id.encode(), int(num_features))
        if rc == -2:
            raise RuntimeError(self.L.l2d_last_error().decode())
        return rc

    def addTemplate_rotate(self, class_id, zero_id, theta, center):
        rc = self.L.l2d_add_template_rotate(self.h, class_id.encode(), int(zero_id), float(theta), float(center[0]), float(center[1]))
        if rc == -2:
            raise RuntimeError(self.L.l2d_last_error().decode())
        return rc

    def numTemplates(self, class_id):
        return self.L.l2d_num_templates(self.h, class_id.encode())

    def export_class(self, class_id, with_theta=False):
        """-> tdesc[n, levels, 5], feats[sum, 3] (and theta[sum]): what Detector.setClassTemplates takes"""
        n, lv, tot = C.c_int(), C.c_int(), C.c_int()
        self._chk(self.L.l2d_class_sizes(self.h, class_id.encode(), C.byref(n), C.byref(lv), C.byref(tot)))
        tdesc = np.zeros((n.value, lv.value, 5), np.int32)
        feats = np.zeros((tot.value, 3), np.int32)
        th = np.zeros(tot.value, np.float32)
        self._chk(self.L.l2d_export_class(self.h, class_id.encode(), tdesc.ctypes.data_as(C.c_void_p), feats.ctypes.data_as(C.c_void_p),
                                          th.ctypes.data_as(C.c_void_p) if with_theta else None))
        return (tdesc, feats, th) if with_theta else (tdesc, feats)

    def readClasses(self, class_ids, fmt):
        return [self._chk(self.L.l2d_read_class(self.h, c.encode(), fmt.encode())) for c in class_ids]

    def writeClasses(self, fmt):
        self._chk(self.L.l2d_write_classes(self.h, fmt.encode()))

    def match(self, img, threshold, class_ids=(), mask=None, cap=1 << 16):
        img, a = _img_args(img)
        m = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            m = mask.ctypes.data_as(C.c_void_p)
        out = np.zeros(cap, MATCH_REC)
        n = C.c_int()
        self._chk(self.L.l2d_match(self.h, *a, m, float(threshold), ",".join(class_ids).encode(), out.ctypes.data_as(C.c_void_p), cap, C.byref(n)))
        assert n.value <= cap
        return out[: n.value].copy()

    @staticmethod
    def groupUniqueId():
        from . import _capi

        _capi.prefer_bundled_nccl()
        buf = C.create_string_buffer(128)
        if lib().l2d_group_unique_id(buf) < 0:
            raise RuntimeError(lib().l2d_last_error().decode())
        return buf.raw

    def setGroup(self, n_ranks, rank, unique_id, mode=0):
        from . import _capi

        _capi.prefer_bundled_nccl()
        self._chk(self.L.l2d_set_group(self.h, int(n_ranks), int(rank), C.c_char_p(unique_id), int(mode)))


def shape_infos(angle_range, scale_range, angle_step=15.0, scale_step=0.5, cap=1 << 16):
    """shapeInfo_producer::produce_infos -> float32 [n, 2] = (angle, scale), scale-major like the reference"""
    L = lib()
    a = np.ascontiguousarray(angle_range, np.float32)
    s = np.ascontiguousarray(scale_range, np.float32)
    out = np.zeros((cap, 2), np.float32)
    n = C.c_int()
    if L.l2d_shape_infos(a.ctypes.data_as(C.c_void_p), len(a), s.ctypes.data_as(C.c_void_p), len(s), angle_step, scale_step,
                         out.ctypes.data_as(C.c_void_p), cap, C.byref(n)) < 0:
        raise RuntimeError(L.l2d_last_error().decode())
    assert n.value <= cap
    return out[: n.value].copy()


def shape_transform(img, angle, scale, as_mask=False):
    """shapeInfo_producer::src_of / mask_of"""
    L = lib()
    img, a = _img_args(img)
    ow, oh = C.c_int(), C.c_int()
    if L.l2d_shape_transform(*a, angle, scale, int(as_mask), None, C.byref(ow), C.byref(oh)) < 0:
        raise RuntimeError(L.l2d_last_error().decode())
    out = np.zeros((oh.value, ow.value, img.shape[2]), np.uint8)
    if L.l2d_shape_transform(*a, angle, scale, int(as_mask), out.ctypes.data_as(C.c_void_p), C.byref(ow), C.byref(oh)) < 0:
        raise RuntimeError(L.l2d_last_error().decode())
    return out[:, :, 0] if out.shape[2] == 1 else out
