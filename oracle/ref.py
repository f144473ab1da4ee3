"""ORACLE / TEST INFRASTRUCTURE -- ctypes front end of oracle/_ref/libsbm_ref.so.

libsbm_ref.so is the reference's own ``line2Dup.cpp`` compiled verbatim (oracle/ref_driver.cpp,
oracle/Makefile).  Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s reference /
cpu_baseline legs may import this module; the product package never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libsbm_ref.so")

MATCH_DTYPE = np.dtype(
    [("x", "<i4"), ("y", "<i4"), ("similarity", "<f4"), ("class_idx", "<i4"), ("template_id", "<i4")]
)


def available():
    return os.path.exists(REF_SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libsbm_ref.so missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int, C.c_float, C.c_float]
        L.ref_destroy.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, a.ctypes.data_as(C.c_void_p)


def _img_args(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, c = img.shape
    return img, (img.ctypes.data_as(C.c_void_p), C.c_int(w), C.c_int(h), C.c_int(c))


class RefDetector:
    """line2Dup::Detector of the reference (/root/reference/line2Dup.h:257-333)."""

    def __init__(self, num_features=63, T=(4, 8), weak=30.0, strong=60.0):
        self.L = lib()
        self.T = list(T)
        arr = (C.c_int * len(T))(*T)
        self.h = C.c_void_p(self.L.ref_create(num_features, arr, len(T), weak, strong))

    def __del__(self):
        try:
            self.L.ref_destroy(self.h)
        except Exception:
            pass

    def _chk(self, rc):
        if rc < 0:
            raise RuntimeError(self.L.ref_last_error().decode())
        return rc

    def read_class(self, class_id, fmt):
        return self._chk(self.L.ref_read_class(self.h, class_id.encode(), fmt.encode()))

    def write_classes(self, fmt):
        return self._chk(self.L.ref_write_classes(self.h, fmt.encode()))

    def add_class(self, class_id, tdesc, feats):
        tdesc = np.ascontiguousarray(tdesc, dtype=np.int32)
        feats = np.ascontiguousarray(feats, dtype=np.int32)
        n, levels = tdesc.shape[0], tdesc.shape[1]
        return self._chk(
            self.L.ref_add_class(
                self.h, class_id.encode(), n, levels, tdesc.ctypes.data_as(C.c_void_p), feats.ctypes.data_as(C.c_void_p)
            )
        )

    def export_class(self, class_id, with_theta=False):
        n, lv, tot = C.c_int(), C.c_int(), C.c_int()
        self._chk(self.L.ref_class_sizes(self.h, class_id.encode(), C.byref(n), C.byref(lv), C.byref(tot)))
        tdesc = np.zeros((n.value, lv.value, 5), np.int32)
        feats = np.zeros((tot.value, 3), np.int32)
        th = np.zeros(tot.value, np.float32)
        self._chk(
            self.L.ref_export_class(
                self.h, class_id.encode(), tdesc.ctypes.data_as(C.c_void_p), feats.ctypes.data_as(C.c_void_p),
                th.ctypes.data_as(C.c_void_p) if with_theta else None,
            )
        )
        return (tdesc, feats, th) if with_theta else (tdesc, feats)

    def add_template(self, img, class_id, mask=None, num_features=0):
        img, a = _img_args(img)
        m = None
        if mask is not None:
            mask, m = _u8(mask)
        return self.L.ref_add_template(self.h, *a, m, class_id.encode(), num_features)

    def add_template_rotate(self, class_id, zero_id, theta, cx, cy):
        return self.L.ref_add_template_rotate(
            self.h, class_id.encode(), zero_id, C.c_float(theta), C.c_float(cx), C.c_float(cy)
        )

    @staticmethod
    def _mask_arg(img, mask):
        """NULL, or the u8 [h][w] mask Detector::match takes as `masks` (line2Dup.cpp:1078-1086)"""
        if mask is None:
            return None, None
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        assert mask.shape == img.shape[:2], "mask.size() != source.size()"
        return mask, mask.ctypes.data_as(C.c_void_p)

    def match(self, img, threshold, class_ids=(), omp_threads=1, cap=1 << 20, mask=None):
        img, a = _img_args(img)
        _keep, m = self._mask_arg(img, mask)
        out = np.zeros(cap, MATCH_DTYPE)
        n = C.c_int()
        self._chk(
            self.L.ref_match(
                self.h, *a, m, C.c_float(threshold), ",".join(class_ids).encode(), omp_threads,
                out.ctypes.data_as(C.c_void_p), cap, C.byref(n),
            )
        )
        return out[: n.value].copy()

    def candidates(self, img, threshold, class_ids=(), cap=1 << 20, mask=None):
        img, a = _img_args(img)
        _keep, m = self._mask_arg(img, mask)
        out = np.zeros(cap, MATCH_DTYPE)
        n = C.c_int()
        self._chk(
            self.L.ref_candidates(
                self.h, *a, m, C.c_float(threshold), ",".join(class_ids).encode(), out.ctypes.data_as(C.c_void_p), cap, C.byref(n)
            )
        )
        return out[: n.value].copy()

    def front(self, img, level, mask=None):
        """-> dict(quantized, spread, response[8], linmem[8, T*T, W*H]) of one pyramid level."""
        img, a = _img_args(img)
        _keep, m = self._mask_arg(img, mask)
        h, w = img.shape[0] >> level, img.shape[1] >> level
        T = self.T[level]
        q = np.zeros((h, w), np.uint8)
        s = np.zeros((h, w), np.uint8)
        r = np.zeros((8, h, w), np.uint8)
        lm = np.zeros((8, T * T, (h // T) * (w // T)), np.uint8)
        self._chk(
            self.L.ref_front(
                self.h, *a, m, level, q.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p),
                r.ctypes.data_as(C.c_void_p), lm.ctypes.data_as(C.c_void_p),
            )
        )
        return dict(quantized=q, spread=s, response=r, linmem=lm)

    def similarity(self, img, class_id, template_id, mask=None):
        img, a = _img_args(img)
        _keep, m = self._mask_arg(img, mask)
        lv = len(self.T) - 1
        T = self.T[lv]
        H, W = (img.shape[0] >> lv) // T, (img.shape[1] >> lv) // T
        out = np.zeros((H, W), np.uint16)
        self._chk(self.L.ref_similarity(self.h, *a, m, class_id.encode(), template_id, out.ctypes.data_as(C.c_void_p)))
        return out

    def similarity_local(self, img, class_id, template_id, level, x, y, mask=None):
        img, a = _img_args(img)
        _keep, m = self._mask_arg(img, mask)
        out = np.zeros((16, 16), np.uint16)
        self._chk(
            self.L.ref_similarity_local(
                self.h, *a, m, class_id.encode(), template_id, level, x, y, out.ctypes.data_as(C.c_void_p)
            )
        )
        return out

    def time(self, img, threshold, class_ids=(), omp_threads=0, iters=3):
        """-> dict(ms_front, ms_match, ms_total, n_matches): mean over iters (see ref_driver.cpp ref_time)."""
        img, a = _img_args(img)
        f, m, t, n = C.c_double(), C.c_double(), C.c_double(), C.c_int()
        self._chk(
            self.L.ref_time(
                self.h, *a, C.c_float(threshold), ",".join(class_ids).encode(), omp_threads, iters,
                C.byref(f), C.byref(m), C.byref(t), C.byref(n),
            )
        )
        return dict(ms_front=f.value, ms_match=m.value, ms_total=t.value, n_matches=n.value)

    def eval_counts(self, img, threshold, class_ids=()):
        """-> (coarse evaluations, coarse candidates, local evaluations) of one frame (ref_driver.cpp ref_eval_counts)."""
        img, a = _img_args(img)
        ce, cc, le = C.c_ulonglong(), C.c_ulonglong(), C.c_ulonglong()
        self._chk(self.L.ref_eval_counts(self.h, *a, C.c_float(threshold), ",".join(class_ids).encode(),
                                         C.byref(ce), C.byref(cc), C.byref(le)))
        return ce.value, cc.value, le.value


def shape_infos(angle_range, scale_range, angle_step=15.0, scale_step=0.5, cap=1 << 16):
    """shapeInfo_producer::produce_infos of the reference (line2Dup.h:407-449) -> float32 [n, 2] = (angle, scale)"""
    L = lib()
    a = np.ascontiguousarray(angle_range, np.float32)
    s = np.ascontiguousarray(scale_range, np.float32)
    out = np.zeros((cap, 2), np.float32)
    n = C.c_int()
    rc = L.ref_shape_infos(a.ctypes.data_as(C.c_void_p), len(a), s.ctypes.data_as(C.c_void_p), len(s), C.c_float(angle_step),
                           C.c_float(scale_step), out.ctypes.data_as(C.c_void_p), cap, C.byref(n))
    if rc < 0:
        raise RuntimeError(L.ref_last_error().decode())
    assert n.value <= cap
    return out[: n.value].copy()


def shape_transform(img, angle, scale, as_mask=False):
    """shapeInfo_producer::src_of / mask_of of the reference (line2Dup.h:379-405, :451-457)"""
    L = lib()
    img, a = _img_args(img)
    ow, oh = C.c_int(), C.c_int()
    rc = L.ref_shape_transform(*a, C.c_float(angle), C.c_float(scale), int(as_mask), None, C.byref(ow), C.byref(oh))
    if rc < 0:
        raise RuntimeError(L.ref_last_error().decode())
    out = np.zeros((oh.value, ow.value, img.shape[2]), np.uint8)
    rc = L.ref_shape_transform(*a, C.c_float(angle), C.c_float(scale), int(as_mask), out.ctypes.data_as(C.c_void_p), C.byref(ow), C.byref(oh))
    if rc < 0:
        raise RuntimeError(L.ref_last_error().decode())
    return out[:, :, 0] if out.shape[2] == 1 else out
