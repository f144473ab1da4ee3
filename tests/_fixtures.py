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


def write_templates_yaml(path, class_id, tdesc, feats):
    """A class in the OpenCV FileStorage YAML 1.0 dialect the reference writes (line2Dup.cpp:86-113, :1548-1575;
    sample: /root/reference/test/case1/test_templ.yaml:1-20)."""
    n, levels = tdesc.shape[:2]
    out = ["%YAML:1.0", "---", "class_id: %s" % class_id, "pyramid_levels: %d" % levels, "template_pyramids:"]
    fo = 0
    for t in range(n):
        out += ["   -", "      template_id: %d" % t, "      templates:"]
        for l in range(levels):
            wd, ht, tx, ty, nf = (int(v) for v in tdesc[t, l])
            out += ["         -", "            width: %d" % wd, "            height: %d" % ht, "            tl_x: %d" % tx,
                    "            tl_y: %d" % ty, "            pyramid_level: %d" % l, "            features:"]
            for x, y, lab in feats[fo : fo + nf]:
                out.append("               - [ %d, %d, %d ]" % (x, y, lab))
            fo += nf
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")


def parse_driver_templates(text):
    """'T t L l w h tlx tly nf n : x,y,label ...' lines of tests/cpp/dropin_driver -> (tdesc[n, levels, 5], feats[sum, 3])"""
    import numpy as np

    rows, feats = {}, []
    for line in text.splitlines():
        if not line.startswith("T "):
            continue
        head, _, tail = line.partition(":")
        h = head.split()
        t, l = int(h[1]), int(h[3])
        rows[(t, l)] = [int(h[4]), int(h[5]), int(h[6]), int(h[7]), int(h[9])]
        feats += [[int(v) for v in tok.split(",")] for tok in tail.split()]
    n = 1 + max(k[0] for k in rows)
    levels = 1 + max(k[1] for k in rows)
    tdesc = np.array([[rows[(t, l)] for l in range(levels)] for t in range(n)], np.int32)
    return tdesc, np.array(feats, np.int32).reshape(-1, 3)


def load_masked_two_class():
    """tests/golden/masked_two_class.npz (make_fixtures.py: make_masked_two_class) -> dict with img, mask and the two classes
    as [(class_id, tdesc, feats)]: "circle" = all case0 templates, "a_ring" = the first `ring_templates` of them"""
    z = dict(np.load(os.path.join(GOLDEN, "masked_two_class.npz")))
    z["img"] = decode_png(z.pop("img_png"))
    z["mask"] = decode_png(z.pop("mask_png"))
    cid, tdesc, feats, nf = load_templates("case0")
    n2 = int(z["ring_templates"])
    z["classes"] = [("circle", tdesc, feats), ("a_ring", tdesc[:n2], feats[: int(tdesc[:n2, :, 4].sum())])]
    z["requests"] = (("all", []), ("circle", ["circle"]), ("ring_circle", ["a_ring", "circle"]), ("circle_ring", ["circle", "a_ring"]))
    return z
