"""Expected values for the synthetic parity tests: the reference's own code (oracle/_ref) wherever the reference
accepts the input, the C++ restatement (oracle/sbm_oracle.cpp) otherwise -- and, where both run, they must agree.

TEST INFRASTRUCTURE: imports oracle/, never imported by the product.

The reference accepts a frame when, at every pyramid level, rows % T == 0 and cols % T == 0 (linearize,
/root/reference/line2Dup.cpp:751-752) and rows * cols is a multiple of the SIMD width (computeResponseMaps, :639 and
:708; mipp::N<uint8_t>() = 32 in the AVX2 build oracle/Makefile makes).  Its refinement clamp (:1231-1236) assumes
the frame is at least 16 T + template wide at every refined level; smaller frames index before the start of a linear
memory (undefined behaviour in the reference), so those stay with the restatement only.
"""
import numpy as np

from oracle import port, ref

REC_KEYS = ("x", "y", "template_id", "class_idx")


def ref_admissible(shape, T, classes):
    h, w = shape[:2]
    L = len(T)
    for l, t in enumerate(T):
        hl, wl = h >> l, w >> l
        if hl <= 0 or wl <= 0 or hl % t or wl % t or (hl * wl) % 32:
            return False
    for _cid, tdesc, _feats in classes:
        tdesc = np.asarray(tdesc)
        if tdesc.shape[0] == 0:
            continue
        for l in range(L - 1):  # refined levels: the clamp window must not be empty
            t = T[l]
            if (w >> l) - int(tdesc[:, l, 0].max()) - 8 * t < 8 * t or (h >> l) - int(tdesc[:, l, 1].max()) - 8 * t < 8 * t:
                return False
        # the coarse scan adds lm[j] for j < template_positions without a bound check: a template larger than the frame
        # gives template_positions <= 0 and is harmless, but a feature row that starts beyond the frame is skipped (:836)
    return True


def same_list(a, b, what=""):
    assert len(a) == len(b), "%s: %d vs %d records" % (what, len(a), len(b))
    for k in REC_KEYS:
        if k in a.dtype.names and k in b.dtype.names:
            assert np.array_equal(a[k], b[k]), (what, k)
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32)), what


class Expected:
    """classes: list of (class_id, tdesc, feats) -- class slots follow the sorted ids (std::map order)."""

    def __init__(self, T, weak, classes, num_features=63, use_ref=True):
        self.T = [int(t) for t in T]
        self.weak = float(weak)
        self.classes = sorted(classes, key=lambda c: c[0])
        self.ids = [c[0] for c in self.classes]
        self.port = port.PortDetector(self.T, self.weak)
        for _cid, td, ft in self.classes:
            self.port.add_class(td, ft)
        self.ref = None
        self._want_ref = use_ref and ref.available()
        self.num_features = num_features
        self.img = self.mask = None

    def process(self, img, mask=None):
        self.img, self.mask = img, mask
        self.port.process(img, mask)
        self.ref = None
        if self._want_ref and ref_admissible(img.shape, self.T, self.classes):
            self.ref = ref.RefDetector(self.num_features, self.T, self.weak)
            for cid, td, ft in self.classes:
                self.ref.add_class(cid, td, ft)
        return self

    @property
    def pinned(self):
        """True when the expectations below come from the reference's own code"""
        return self.ref is not None

    def intermediate(self, level, kind):
        """kind 0..3 = quantized, spread, response, linear memories (flat bytes)"""
        a = self.port.get(level, kind)
        if self.ref is not None:
            f = self.ref.front(self.img, level, mask=self.mask)
            b = (f["quantized"], f["spread"], f["response"], f["linmem"])[kind].ravel()
            assert np.array_equal(a, b), "restatement and reference disagree at level %d kind %d" % (level, kind)
            return b
        return a

    def similarity(self, class_id, template_id):
        a = self.port.similarity(self.ids.index(class_id), template_id)
        if self.ref is not None:
            b = self.ref.similarity(self.img, class_id, template_id, mask=self.mask)
            assert np.array_equal(a, b), "restatement and reference disagree on the raw map of %s/%d" % (class_id, template_id)
            return b
        return a

    def similarity_local(self, class_id, template_id, level, x, y):
        a = self.port.similarity_local(self.ids.index(class_id), template_id, level, x, y)
        if self.ref is not None:
            b = self.ref.similarity_local(self.img, class_id, template_id, level, x, y, mask=self.mask)
            assert np.array_equal(a, b), "restatement and reference disagree on a local patch"
            return b
        return a

    def match(self, threshold, class_ids=(), finalize=True):
        req = list(class_ids) if class_ids else list(self.ids)
        idx = [self.ids.index(c) for c in req if c in self.ids]
        a = self.port.match(threshold, class_idx=tuple(idx), finalize=finalize) if idx else np.zeros(0, port.REC)
        if self.ref is not None:
            if not idx:
                return a
            if finalize:
                b = self.ref.match(self.img, threshold, req, omp_threads=1, mask=self.mask)
            else:
                b = self.ref.candidates(self.img, threshold, req, mask=self.mask)
            same_list(a, b, "restatement vs reference (%s)" % ("final" if finalize else "candidates"))
            return b
        return a
