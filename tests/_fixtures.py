"""Loaders for the committed fixtures under tests/golden/ (written by tests/golden/make_fixtures.py)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

CASES = ["case1", "case2", "case0_1", "case0_2", "case0_3", "m1"]


def decode_png(buf):
    import cv2

    img = cv2.imdecode(np.asarray(buf, np.uint8), cv2.IMREAD_UNCHANGED)
    assert img is not None
    return img


def load_templates(name):
    """-> class_id, tdesc[n, levels, 5] int32, feats[sum, 3] int32, num_features"""
    limit = None
    if name.endswith("]"):  # "case1[:360]"
        name, sl = name[:-1].split("[:")
        limit = int(sl)
    z = np.load(os.path.join(GOLDEN, "templates_%s.npz" % name))
    tdesc = z["tdesc"].astype(np.int32)
    feats = z["feats"].astype(np.int32)
    if limit is not None:
        nf = int(tdesc[:limit, :, 4].sum())
        tdesc, feats = tdesc[:limit], feats[:nf]
    return str(z["class_id"]), tdesc, feats, int(z["num_features"])


def load_case(name):
    z = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    z["img"] = decode_png(z.pop("img_png"))
    assert tuple(z["img"].shape) == tuple(z["shape"])
    z["templates"] = str(z["templates"])
    z["class_id"] = str(z["class_id"])
    return z


def to_gray(img):
    import cv2

    return cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
