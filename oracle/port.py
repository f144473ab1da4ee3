"""ORACLE / TEST INFRASTRUCTURE -- ctypes front end of oracle/libsbm_oracle.so (oracle/sbm_oracle.cpp, the plain C++
restatement of the match path).  Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import it."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "libsbm_oracle.so")
REC = np.dtype([("x", "<i4"), ("y", "<i4"), ("similarity", "<f4"), ("class_idx", "<i4"), ("template_id", "<i4")])
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            subprocess.check_call(["make", "-C", _HERE, "port"])
        L = C.CDLL(SO)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_float]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_add_class.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_get.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_similarity.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_similarity_local.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_match.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        _lib = L
    return _lib


class PortDetector:
    def __init__(self, T=(4, 8), weak=30.0):
        self.L = lib()
        self.T = [int(t) for t in T]
        arr = (C.c_int * len(self.T))(*self.T)
        self.h = C.c_void_p(self.L.orc_create(arr, len(self.T), weak))
        self.shape = None

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    def add_class(self, tdesc, feats):
        tdesc = np.ascontiguousarray(tdesc, np.int32)
        feats = np.ascontiguousarray(feats, np.int32)
        return self.L.orc_add_class(self.h, tdesc.shape[0], tdesc.ctypes.data_as(C.c_void_p), feats.ctypes.data_as(C.c_void_p))

    def process(self, img, mask=None):
        img = np.ascontiguousarray(img, np.uint8)
        if img.ndim == 2:
            img = img[:, :, None]
        h, w, c = img.shape
        m = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
            m = mask.ctypes.data_as(C.c_void_p)
        self.L.orc_process(self.h, img.ctypes.data_as(C.c_void_p), w, h, c, m)
        self.shape = (h, w)

    def get(self, level, kind):
        h, w = self.shape[0] >> level, self.shape[1] >> level
        n = h * w * (1 if kind <= 1 else 8)
        out = np.zeros(n, np.uint8)
        self.L.orc_get(self.h, level, kind, out.ctypes.data_as(C.c_void_p))
        return out

    def similarity(self, class_idx, template_id):
        lv = len(self.T) - 1
        H, W = (self.shape[0] >> lv) // self.T[lv], (self.shape[1] >> lv) // self.T[lv]
        out = np.zeros((H, W), np.uint16)
        self.L.orc_similarity(self.h, class_idx, template_id, out.ctypes.data_as(C.c_void_p))
        return out

    def similarity_local(self, class_idx, template_id, level, x, y):
        out = np.zeros((16, 16), np.uint16)
        self.L.orc_similarity_local(self.h, class_idx, template_id, level, x, y, out.ctypes.data_as(C.c_void_p))
        return out

    def match(self, threshold, class_idx=(0,), finalize=True, cap=1 << 20):
        arr = (C.c_int * len(class_idx))(*class_idx)
        out = np.zeros(cap, REC)
        n = self.L.orc_match(self.h, threshold, arr, len(class_idx), int(finalize), out.ctypes.data_as(C.c_void_p), cap)
        assert n <= cap
        return out[:n].copy()
