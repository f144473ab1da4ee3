#!/usr/bin/env python
"""Generate the committed parity fixtures under tests/golden/ (run in the BUILD container only).

Needs /root/reference (fixture images + template YAMLs), Python cv2 4.13.0 (the executable
specification of the OpenCV ops the reference calls) and oracle/_ref/libsbm_ref.so (the reference's
own line2Dup.cpp compiled verbatim: `make -C oracle ref`).  Nothing under tests/ reads
/root/reference at run time: every input and every expected output the tests use is written here.

Outputs (all .npz, zlib-compressed; images stored as PNG bytes, decoded with cv2.imdecode):
  cv_pins.npz            random inputs + REAL cv2 outputs for GaussianBlur/Sobel/phase/convert/pyrDown/...
  templates_<set>.npz    tdesc[n,levels,5]={width,height,tl_x,tl_y,nf}, feats[sum,3]={x,y,label}
  <case>.npz             frame (prepared exactly like the reference's test drivers do), detector
                         parameters, expected outputs of the reference (final list at 1 OpenMP thread,
                         pre-unique candidate list, quantised images from the cv2 pipeline AND _ref,
                         sha256 of spread/response/linear memories, a few raw similarity maps)
"""
import hashlib
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.ref import RefDetector  # noqa: E402

REF = "/root/reference/test/"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def png(img):
    ok, buf = cv2.imencode(".png", img, [cv2.IMWRITE_PNG_COMPRESSION, 9])
    assert ok
    return np.frombuffer(buf.tobytes(), np.uint8)


# ------------------------------------------------------------------------------------------------
# numpy + REAL cv2 restatement of quantizedOrientations + hysteresisGradient (line2Dup.cpp:218-404),
# used only here, to tie the goldens to real OpenCV arithmetic rather than to oracle/shim.
# ------------------------------------------------------------------------------------------------
def cv2_quantize(img, weak=30.0):
    sm = cv2.GaussianBlur(img, (7, 7), 0, borderType=cv2.BORDER_REPLICATE)
    if img.ndim == 2 or img.shape[2] == 1:
        dx = cv2.Sobel(sm, cv2.CV_32F, 1, 0, ksize=3, borderType=cv2.BORDER_REPLICATE)
        dy = cv2.Sobel(sm, cv2.CV_32F, 0, 1, ksize=3, borderType=cv2.BORDER_REPLICATE)
        mag = dx * dx + dy * dy
    else:
        dx3 = cv2.Sobel(sm, cv2.CV_16S, 1, 0, ksize=3, borderType=cv2.BORDER_REPLICATE).astype(np.int32)
        dy3 = cv2.Sobel(sm, cv2.CV_16S, 0, 1, ksize=3, borderType=cv2.BORDER_REPLICATE).astype(np.int32)
        m3 = dx3 * dx3 + dy3 * dy3
        m1, m2, m3_ = m3[..., 0], m3[..., 1], m3[..., 2]
        pick = np.where((m1 >= m2) & (m1 >= m3_), 0, np.where((m2 >= m1) & (m2 >= m3_), 1, 2))
        dx = np.take_along_axis(dx3, pick[..., None], 2)[..., 0].astype(np.float32)
        dy = np.take_along_axis(dy3, pick[..., None], 2)[..., 0].astype(np.float32)
        mag = np.take_along_axis(m3, pick[..., None], 2)[..., 0].astype(np.float32)
    ang = cv2.phase(dx, dy, angleInDegrees=True)
    q = cv2.convertScaleAbs(ang, alpha=16.0 / 360.0)
    q[0, :] = 0
    q[-1, :] = 0
    q[:, 0] = 0
    q[:, -1] = 0
    q[1:-1, 1:-1] &= 7
    h, w = q.shape
    hist = np.zeros((8, h - 2, w - 2), np.int32)
    for dy_ in range(3):
        for dx_ in range(3):
            nb = q[dy_ : dy_ + h - 2, dx_ : dx_ + w - 2]
            for b in range(8):
                hist[b] += nb == b
    idx = hist.argmax(0)  # first maximum == lowest bin among ties
    votes = hist.max(0)
    out = np.zeros((h, w), np.uint8)
    ok = (mag[1:-1, 1:-1] > np.float32(weak * weak)) & (votes >= 5)
    out[1:-1, 1:-1] = np.where(ok, (1 << idx).astype(np.uint8), 0)
    return out


def cv2_pyramid_quantized(img, levels, weak=30.0):
    out = []
    src = img
    for l in range(levels):
        if l > 0:
            src = cv2.pyrDown(src, dstsize=(src.shape[1] // 2, src.shape[0] // 2))
        out.append(cv2_quantize(src, weak))
    return out


# ------------------------------------------------------------------------------------------------
def make_cv_pins():
    rng = np.random.default_rng(7)
    d = {}
    g = rng.integers(0, 256, (96, 128), dtype=np.uint8)
    c = rng.integers(0, 256, (64, 80, 3), dtype=np.uint8)
    # smooth-ish content as well (noise alone under-exercises rounding ties)
    s = cv2.GaussianBlur(rng.integers(0, 256, (72, 88, 3), dtype=np.uint8), (9, 9), 3)
    for name, im in (("g", g), ("c", c), ("s", s)):
        d[name] = im
        d[name + "_gauss"] = cv2.GaussianBlur(im, (7, 7), 0, borderType=cv2.BORDER_REPLICATE)
        d[name + "_sobx"] = cv2.Sobel(im, cv2.CV_16S, 1, 0, ksize=3, borderType=cv2.BORDER_REPLICATE)
        d[name + "_soby"] = cv2.Sobel(im, cv2.CV_16S, 0, 1, ksize=3, borderType=cv2.BORDER_REPLICATE)
        d[name + "_pyr"] = cv2.pyrDown(im, dstsize=(im.shape[1] // 2, im.shape[0] // 2))
    odd = rng.integers(0, 256, (50, 70), dtype=np.uint8)
    d["odd"] = odd
    d["odd_pyr"] = cv2.pyrDown(odd, dstsize=(35, 25))
    # phase: every integer pair the Sobel can produce is in [-1020, 1020]^2; sample 2^18 + axes/diagonals
    px = rng.integers(-1020, 1021, 1 << 18).astype(np.float32)
    py = rng.integers(-1020, 1021, 1 << 18).astype(np.float32)
    ax = np.arange(-1020, 1021, dtype=np.float32)
    px = np.concatenate([px, ax, np.zeros_like(ax), ax, ax, [0]]).astype(np.float32)
    py = np.concatenate([py, np.zeros_like(ax), ax, ax, -ax, [0]]).astype(np.float32)
    d["phase_x"], d["phase_y"] = px, py
    d["phase_deg"] = cv2.phase(px.reshape(1, -1), py.reshape(1, -1), angleInDegrees=True).ravel()
    # convertTo(CV_8U, 16/360): rint half-to-even on float product
    a = np.concatenate([rng.uniform(0, 360, 1 << 16).astype(np.float32),
                        (np.arange(0, 33, dtype=np.float32) * np.float32(11.25)),
                        np.float32([0, 359.99997, 360.0, 11.249999, 11.25, 11.250001, 33.75, 348.75])])
    d["conv_in"] = a
    d["conv_out"] = cv2.convertScaleAbs(a.reshape(1, -1), alpha=16.0 / 360.0).ravel()
    # erode 3x3 replicate (training path)
    m = (rng.random((40, 56)) > 0.2).astype(np.uint8) * 255
    d["mask"] = m
    d["mask_erode"] = cv2.erode(m, None, borderType=cv2.BORDER_REPLICATE)
    # bilinear resize (training path, shapeInfo_producer::transform)
    for i, f in enumerate((0.5, 0.8, 1.1, 0.37)):
        d["rs%d_f" % i] = np.float64(f)
        d["rs%d" % i] = cv2.resize(c, None, fx=f, fy=f)
        d["rsg%d" % i] = cv2.resize(g, None, fx=f, fy=f)
    np.savez_compressed(os.path.join(HERE, "cv_pins.npz"), **d)
    print("cv_pins.npz written")


def export_templates(name, class_id, fmt, num_features, limit=None):
    det = RefDetector(num_features, (4, 8))
    n = det.read_class(class_id, fmt)
    tdesc, feats = det.export_class(class_id)
    if limit is not None and limit < n:
        nf = tdesc[:limit, :, 4].sum()
        tdesc, feats = tdesc[:limit], feats[:nf]
    np.savez_compressed(
        os.path.join(HERE, "templates_%s.npz" % name),
        class_id=np.array(class_id), tdesc=tdesc.astype(np.int32), feats=feats.astype(np.int16),
        num_features=np.int32(num_features), T=np.int32([4, 8]),
    )
    print("templates_%s.npz: %d templates, %d features" % (name, tdesc.shape[0], feats.shape[0]))
    return tdesc, feats


def make_case(name, img, tset, class_id, tdesc, feats, num_features, sim_ids, thr=90.0):
    det = RefDetector(num_features, (4, 8))
    det.add_class(class_id, tdesc, feats)
    d = dict(
        img_png=png(img), shape=np.int32(img.shape), templates=np.array(tset), class_id=np.array(class_id),
        num_features=np.int32(num_features), T=np.int32([4, 8]), weak=np.float32(30), strong=np.float32(60),
        threshold=np.float32(thr),
    )
    d["matches"] = det.match(img, thr, [class_id], omp_threads=1)
    d["candidates"] = det.candidates(img, thr, [class_id])
    qcv = cv2_pyramid_quantized(img, 2)
    for l in range(2):
        f = det.front(img, l)
        assert np.array_equal(f["quantized"], qcv[l]), "oracle/shim disagrees with real cv2 at level %d of %s" % (l, name)
        d["quantized%d" % l] = f["quantized"]
        d["spread%d_sha" % l] = np.array(sha(f["spread"]))
        d["response%d_sha" % l] = np.array(sha(f["response"]))
        d["linmem%d_sha" % l] = np.array(sha(f["linmem"]))
        # a thin slice of the linear memories in the clear, to localise a mismatch
        d["linmem%d_label3" % l] = f["linmem"][3]
    d["sim_ids"] = np.int32(sim_ids)
    d["sim_maps"] = np.stack([det.similarity(img, class_id, t) for t in sim_ids])
    # gray variant of the same frame exercises the 1-channel branch (line2Dup.cpp:322-330)
    gray = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    d["gray_matches"] = det.match(gray, thr, [class_id], omp_threads=1)
    d["gray_candidates"] = det.candidates(gray, thr, [class_id])
    qg = cv2_pyramid_quantized(gray, 2)
    for l in range(2):
        f = det.front(gray, l)
        assert np.array_equal(f["quantized"], qg[l])
        d["gray_quantized%d" % l] = f["quantized"]
        d["gray_linmem%d_sha" % l] = np.array(sha(f["linmem"]))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
    print("%s.npz: frame %s, %d final matches (%d pre-unique); gray %d (%d)" % (
        name, img.shape, len(d["matches"]), len(d["candidates"]), len(d["gray_matches"]), len(d["gray_candidates"])))
    if len(d["matches"]):
        print("   best:", d["matches"][0])


def crop_to(img, stride):
    h, w = img.shape[:2]
    return np.ascontiguousarray(img[: h // stride * stride, : w // stride * stride])


def synth_m1(seed=1234):
    """1024x1024x3, smooth low-contrast background + 4 rotated copies of the case1 object."""
    rng = np.random.default_rng(seed)
    bg = cv2.GaussianBlur(rng.integers(0, 256, (1024, 1024, 3), dtype=np.uint8), (5, 5), 0)
    frame = (bg.astype(np.float32) * 0.25).astype(np.uint8)
    train = cv2.imread(REF + "case1/train.png")
    obj = train[110:380, 130:400]
    pad = 60
    big = np.zeros((obj.shape[0] + 2 * pad, obj.shape[1] + 2 * pad, 3), np.uint8)
    big[pad:-pad, pad:-pad] = obj
    centres = [(260, 270), (760, 250), (270, 760), (740, 760)]
    for k, (cx, cy) in enumerate(centres):
        ang = (k + 1) * 37.0
        M = cv2.getRotationMatrix2D((big.shape[1] / 2.0, big.shape[0] / 2.0), ang, 1.0)
        rot = cv2.warpAffine(big, M, (big.shape[1], big.shape[0]))
        msk = cv2.warpAffine(np.full(big.shape[:2], 255, np.uint8), M, (big.shape[1], big.shape[0])) > 0
        inner = cv2.warpAffine(np.pad(np.full(obj.shape[:2], 255, np.uint8), pad), M, (big.shape[1], big.shape[0])) > 127
        y0, x0 = cy - big.shape[0] // 2, cx - big.shape[1] // 2
        roi = frame[y0 : y0 + big.shape[0], x0 : x0 + big.shape[1]]
        roi[inner] = rot[inner]
        del msk
    return frame


def make_masked_two_class():
    """masked_two_class.npz: Detector::match with a mask (line2Dup.cpp:1078-1086, :439, :446-450) and with two classes in
    the caller's order (:1124-1140), both through the reference's own code.  Frame: test/case0/2.jpg (640x480); classes
    "circle" (the 89 case0 templates) and "a_ring" (case0 templates 0..59 under a name that sorts first, so the class
    slots (std::map order) differ from the insertion order and both classes report the same places)."""
    td0, ft0 = export_templates("case0", "circle", REF + "case0/%s_templ.yaml", 150)
    n2 = 60
    td2, ft2 = td0[:n2], ft0[: td0[:n2, :, 4].sum()]
    img = crop_to(cv2.imread(REF + "case0/2.jpg"), 32)
    h, w = img.shape[:2]
    mask = np.full((h, w), 255, np.uint8)
    mask[:, : w // 3] = 0          # a third of the frame removed
    mask[h // 2 :, ::5] = 0        # and a comb over the lower half (any non-zero value keeps a pixel)
    mask[5:40, w // 2 : w // 2 + 60] = 3
    det = RefDetector(150, (4, 8))
    det.add_class("circle", td0, ft0)
    det.add_class("a_ring", td2, ft2)
    d = dict(img_png=png(img), shape=np.int32(img.shape), mask_png=png(mask), T=np.int32([4, 8]), weak=np.float32(30),
             strong=np.float32(60), num_features=np.int32(150), ring_templates=np.int32(n2))
    thr = 70.0
    d["threshold"] = np.float32(thr)
    for tag, ids in (("all", []), ("circle", ["circle"]), ("ring_circle", ["a_ring", "circle"]), ("circle_ring", ["circle", "a_ring"])):
        d["matches_" + tag] = det.match(img, thr, ids, omp_threads=1)
        d["candidates_" + tag] = det.candidates(img, thr, ids)
        d["masked_matches_" + tag] = det.match(img, thr, ids, omp_threads=1, mask=mask)
        d["masked_candidates_" + tag] = det.candidates(img, thr, ids, mask=mask)
        print("masked_two_class[%s]: %d final (%d pre-unique); masked %d (%d)" % (
            tag, len(d["matches_" + tag]), len(d["candidates_" + tag]), len(d["masked_matches_" + tag]), len(d["masked_candidates_" + tag])))
    assert set(d["matches_all"]["class_idx"]) == {0, 1} and set(d["masked_matches_all"]["class_idx"]) == {0, 1}
    for l in range(2):
        f = det.front(img, l, mask=mask)
        d["masked_quantized%d" % l] = f["quantized"]
        d["masked_linmem%d_sha" % l] = np.array(sha(f["linmem"]))
    d["masked_sim_ring7"] = det.similarity(img, "a_ring", 7, mask=mask)
    np.savez_compressed(os.path.join(HERE, "masked_two_class.npz"), **d)


def main():
    os.makedirs(HERE, exist_ok=True)
    if "--masked-two-class" in sys.argv:  # only this fixture (the others are unchanged)
        make_masked_two_class()
        return
    make_cv_pins()

    # ---- case1: test.cpp angle_test (test.cpp:262-420) ----
    td1, ft1 = export_templates("case1", "test", REF + "case1/%s_templ.yaml", 128)
    im = cv2.imread(REF + "case1/test.png")
    p = 250
    padded = np.zeros((im.shape[0] + 2 * p, im.shape[1] + 2 * p, 3), np.uint8)
    padded[p : p + im.shape[0], p : p + im.shape[1]] = im
    make_case("case1", crop_to(padded, 16), "case1", "test", td1, ft1, 128, [0, 90, 340, 341])

    # ---- case2: test.cpp noise_test (test.cpp:422-524) ----
    td2, ft2 = export_templates("case2", "test", REF + "case2/%s_templ.yaml", 30)
    make_case("case2", crop_to(cv2.imread(REF + "case2/test.png"), 16), "case2", "test", td2, ft2, 30, [0, 100, 200])

    # ---- case0: test.cpp scale_test (test.cpp:162-260) ----
    td0, ft0 = export_templates("case0", "circle", REF + "case0/%s_templ.yaml", 150)
    make_case("case0_1", crop_to(cv2.imread(REF + "case0/1.jpg"), 32), "case0", "circle", td0, ft0, 150, [0, 3, 40, 88])
    make_case("case0_2", crop_to(cv2.imread(REF + "case0/2.jpg"), 32), "case0", "circle", td0, ft0, 150, [0, 88])
    make_case("case0_3", crop_to(cv2.imread(REF + "case0/3.png"), 32), "case0", "circle", td0, ft0, 150, [12, 88])

    # ---- M1 headline: synthetic 1024^2 x case1 templates 0..359 ----
    # templates: the first 360 of templates_case1.npz (the tests slice them the same way)
    n = 360
    nf = td1[:n, :, 4].sum()
    make_case("m1", synth_m1(), "case1[:360]", "test", td1[:n], ft1[:nf], 128, [0, 37, 74, 111, 148])
    make_masked_two_class()


if __name__ == "__main__":
    main()
