"""Post-match NMS: include/nms.hpp against the reference's own nms.hpp (compiled here against
oracle/shim when /root/reference exists) and against a numpy restatement that travels everywhere."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "nms_driver.cpp")
CXX = "/usr/bin/g++"


def build(tmp_path, reference):
    exe = str(tmp_path / ("nms_ref" if reference else "nms_ours"))
    cmd = [CXX, "-std=c++14", "-O2", "-w", SRC, "-o", exe, "-I" + os.path.join(ROOT, "include")]
    if reference:
        cmd += ["-DSBM_NMS_REFERENCE", "-I/root/reference", "-I" + os.path.join(ROOT, "oracle", "shim")]
    else:
        cmd += ["-DSBM_FORCE_CV_COMPAT"]
    subprocess.check_call(cmd)
    return exe


def run(exe, boxes, scores, st, nt, eta, top_k):
    lines = ["%d %r %r %r %d" % (len(boxes), st, nt, eta, top_k)]
    lines += ["%d %d %d %d %r" % (b[0], b[1], b[2], b[3], float(s)) for b, s in zip(boxes, scores)]
    out = subprocess.run([exe], input="\n".join(lines) + "\n", stdout=subprocess.PIPE, text=True, check=True).stdout
    return [int(v) for v in out.split()]


def nms_numpy(boxes, scores, st, nt, eta, top_k):
    """/root/reference/nms.hpp:21-95 restated: float32 scores, IoU = 1.f - (float)(1.0 - inter / union)"""
    scores = np.asarray(scores, np.float32)
    idx = [i for i in range(len(scores)) if scores[i] > np.float32(st)]
    idx.sort(key=lambda i: -float(scores[i]))  # Python's sort is stable
    if 0 < top_k < len(idx):
        idx = idx[:top_k]
    keep, limit = [], np.float32(nt)
    for i in idx:
        ok = True
        for k in keep:
            a, b = boxes[i], boxes[k]
            aa, ab = a[2] * a[3], b[2] * b[3]
            if aa + ab <= 0:
                ov = np.float32(1.0)
            else:
                iw = min(a[0] + a[2], b[0] + b[2]) - max(a[0], b[0])
                ih = min(a[1] + a[3], b[1] + b[3]) - max(a[1], b[1])
                inter = float(iw * ih) if iw > 0 and ih > 0 else 0.0
                ov = np.float32(1.0) - np.float32(1.0 - inter / (aa + ab - inter))
            if not ov <= limit:
                ok = False
                break
        if ok:
            keep.append(i)
            if eta < 1 and limit > 0.5:
                limit = np.float32(limit * np.float32(eta))
    return keep


def cases():
    rng = np.random.default_rng(11)
    out = []
    for n, st, nt, eta, top_k in [(200, 0.0, 0.5, 1.0, 0), (300, 50.0, 0.3, 1.0, 0), (300, 0.0, 0.8, 0.9, 0), (250, 10.0, 0.6, 0.95, 40),
                                  (1, 0.0, 0.5, 1.0, 0), (0, 0.0, 0.5, 1.0, 0)]:
        xy = rng.integers(0, 400, (n, 2))
        wh = rng.integers(0, 120, (n, 2))  # includes empty boxes
        boxes = np.concatenate([xy, wh], 1).tolist()
        scores = np.round(rng.uniform(0, 100, n), 1).astype(np.float32)  # rounded: ties occur, stable order matters
        out.append((boxes, scores, st, nt, eta, top_k))
    return out


def test_nms_equals_restatement(tmp_path):
    exe = build(tmp_path, reference=False)
    for c in cases():
        assert run(exe, *c) == nms_numpy(*c)


@pytest.mark.skipif(not os.path.exists("/root/reference/nms.hpp"), reason="reference tree not present")
def test_nms_equals_reference(tmp_path):
    ours, ref = build(tmp_path, reference=False), build(tmp_path, reference=True)
    for c in cases():
        assert run(ours, *c) == run(ref, *c)
