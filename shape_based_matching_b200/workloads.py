"""Synthetic workloads of BASELINE.md section 2 (M4 = BASELINE.json config 4, M5 = config 5), seeded and reproducible.

Used by bench.py, tools/ and tests/: frames are drawn with cv2 (polygons / ellipses: dense, coherent edges), template
sets are PRODUCED the way the reference's callers produce them (/root/reference/test.cpp:262-330): a base template
trained by Detector::addTemplate on shapeInfo_producer::src_of(info) and expanded with Detector::addTemplate_rotate
over shapeInfo_producer::infos -- through the drop-in C++ classes (line2dup.CxxDetector), whose front end runs on the
GPU.  `trainer` lets a test substitute the reference's own Detector (oracle/ref.RefDetector, same calls).
"""
import numpy as np


def dense_image(rng, h, w, c, n_shapes, max_r=90):
    import cv2

    img = np.full((h, w, c), 40, np.uint8)
    for _ in range(n_shapes):
        col = [int(v) for v in rng.integers(0, 256, c)]
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        if rng.random() < 0.5:
            pts = (np.array([cx, cy]) + rng.integers(-max_r, max_r + 1, (5, 2))).astype(np.int32)
            cv2.fillPoly(img, [pts], col)
        else:
            cv2.ellipse(img, (cx, cy), (int(rng.integers(10, max_r)), int(rng.integers(10, max_r))), float(rng.integers(0, 180)), 0, 360, col, -1)
    return np.ascontiguousarray(img.reshape(h, w, c))


def textured_object(seed, size, n_shapes, max_r):
    """a square gray object full of edges, on a black pad wide enough to rotate it without clipping (test.cpp:273-279)"""
    rng = np.random.default_rng(seed)
    obj = dense_image(rng, size, size, 1, n_shapes, max_r)[:, :, 0]
    pad = int(np.ceil(size * (np.sqrt(2.0) - 1.0) / 2.0)) + 8
    padded = np.zeros((size + 2 * pad, size + 2 * pad), np.uint8)
    padded[pad : pad + size, pad : pad + size] = obj
    mask = np.zeros_like(padded)
    mask[pad : pad + size, pad : pad + size] = 255
    return padded, mask


def _make_trainer(num_features, T, trainer, device):
    if trainer is not None:
        return trainer
    from . import line2dup

    return line2dup.CxxDetector(num_features, T, 30.0, 60.0, device=device)


def rotation_scale_templates(obj, mask, class_id, num_features, angle_range, angle_step, scale_range, scale_step, T=(4, 8), trainer=None,
                             device=0, infos=None, probe=None):
    """test.cpp:281-316 with use_rot: per scale, addTemplate on src_of(first info of that scale), then addTemplate_rotate for
    every further angle of that scale about the centre of the (scaled) image.  -> (tdesc, feats, infos[n, 2]).

    selectScatteredFeatures (line2Dup.cpp:163-212) stops at the first pass that yields AT LEAST num_features, so a request of
    8191 can return more than 8191 features -- which Detector::match then refuses (CV_Assert(features.size() < 8192),
    :811, :1195).  probe(): a factory of throw-away trainers; when given, a base is first trained there and the request is
    lowered by 5 % at a time until every level stays under 8192; the accepted request is what the real class is trained with."""
    from . import line2dup

    det = _make_trainer(num_features, T, trainer, device)
    if infos is None:
        infos = line2dup.shape_infos(angle_range, scale_range, angle_step, scale_step)
    add = getattr(det, "addTemplate", None) or det.add_template
    rot = getattr(det, "addTemplate_rotate", None) or (lambda cid, z, th, c: det.add_template_rotate(cid, z, th, c[0], c[1]))
    first_id, first_angle, cur_scale, centre = None, 0.0, None, None
    kept = []
    for angle, scale in infos:
        if cur_scale is None or scale != cur_scale:  # a new scale: re-extract the features (test.cpp:292-293)
            src = line2dup.shape_transform(obj, float(angle), float(scale))
            msk = line2dup.shape_transform(mask, float(angle), float(scale), as_mask=True)
            want = num_features
            while probe is not None:
                p = probe()
                p_add = getattr(p, "addTemplate", None) or p.add_template
                if p_add(src, "probe", msk, want) < 0:
                    raise RuntimeError("addTemplate found too few features at scale %g" % scale)
                if int(p.export_class("probe")[0][:, :, 4].max()) < 8192:
                    break
                want = int(want * 0.95)
            tid = add(src, class_id, msk, want)
            if tid < 0:
                raise RuntimeError("addTemplate found too few features at scale %g" % scale)
            first_id, first_angle, cur_scale = tid, float(angle), scale
            centre = (src.shape[1] / 2.0, src.shape[0] / 2.0)
        else:
            tid = rot(class_id, first_id, float(angle) - first_angle, centre)
        kept.append((angle, scale))
    tdesc, feats = det.export_class(class_id)
    if trainer is None:
        det.close()
    return tdesc, feats, np.array(kept, np.float32)


def paste_rotated(frame, obj, mask, centre, angle_deg, scale):
    """the object, turned by angle_deg (counter-clockwise on screen: the sense of addTemplate_rotate's theta, whose features
    are turned by -theta in y-down coordinates, line2Dup.cpp:1395-1409) and scaled, pasted at centre"""
    import cv2

    h, w = obj.shape[:2]
    M = cv2.getRotationMatrix2D((w / 2.0, h / 2.0), angle_deg, scale)
    M[0, 2] += centre[0] - w / 2.0
    M[1, 2] += centre[1] - h / 2.0
    warped = cv2.warpAffine(obj, M, (frame.shape[1], frame.shape[0]), flags=cv2.INTER_LINEAR)
    inner = cv2.warpAffine(mask, M, (frame.shape[1], frame.shape[0]), flags=cv2.INTER_NEAREST) > 127
    view = frame if frame.ndim == 2 else frame[:, :, 0]
    view[inner] = warped[inner]


# ---------------------------------------------------------------------------------------------------------------------
# M4 = BASELINE.json config 4: 4096 x 4096 x 1 frame (seed 4096) x 3600 templates (0.1 degree x 3 scales), up to
# 8191 / 4095 features, T = {4, 8}, threshold 90
# ---------------------------------------------------------------------------------------------------------------------
M4 = dict(hw=4096, seed=4096, num_features=8191, angle_range=(0.0, 119.9), angle_step=0.1, scale_range=(0.9, 1.1), scale_step=0.1,
          threshold=90.0, object_size=1100, object_shapes=9000, object_max_r=24, frame_shapes=5000, class_id="m4")


def m4_templates(trainer=None, device=0, n_angles=None, probe=None):
    """-> (tdesc[3600, 2, 5], feats, infos).  n_angles: only the first n_angles angles of every scale (tests)."""
    from . import line2dup

    if probe is None:
        probe = lambda: line2dup.CxxDetector(M4["num_features"], (4, 8), 30.0, 60.0, device=device)  # noqa: E731

    obj, mask = textured_object(M4["seed"] + 1, M4["object_size"], M4["object_shapes"], M4["object_max_r"])
    infos = line2dup.shape_infos(M4["angle_range"], M4["scale_range"], M4["angle_step"], M4["scale_step"])
    if n_angles is not None:
        keep = []
        for s in np.unique(infos[:, 1]):
            keep.append(infos[infos[:, 1] == s][:n_angles])
        infos = np.concatenate(keep)
    return rotation_scale_templates(obj, mask, M4["class_id"], M4["num_features"], None, None, None, None, trainer=trainer, device=device,
                                    infos=infos, probe=probe)


def m4_frame():
    rng = np.random.default_rng(M4["seed"])
    hw = M4["hw"]
    frame = dense_image(rng, hw, hw, 1, M4["frame_shapes"])[:, :, 0]
    obj, mask = textured_object(M4["seed"] + 1, M4["object_size"], M4["object_shapes"], M4["object_max_r"])
    # three true instances: angles on the 0.1-degree grid, one per trained scale
    for centre, angle, scale in (((1000, 1050), 12.3, 1.0), ((3050, 1000), 77.7, 0.9), ((2050, 3050), 101.0, 1.1)):
        paste_rotated(frame, obj, mask, centre, angle, scale)
    return np.ascontiguousarray(frame)


# ---------------------------------------------------------------------------------------------------------------------
# M5 = BASELINE.json config 5: 256 frames of 2048 x 2048 x 1 (seeds 0 .. 255) x 720 templates (0.5 degree), 128 / 64
# features, T = {4, 8}, threshold 90
# ---------------------------------------------------------------------------------------------------------------------
M5 = dict(hw=2048, n_frames=256, num_features=128, angle_range=(0.0, 359.9), angle_step=0.5, threshold=90.0, object_size=240,
          object_shapes=60, object_max_r=45, frame_shapes=700, class_id="m5")


def m5_templates(trainer=None, device=0, n_angles=None):
    from . import line2dup

    obj, mask = textured_object(5, M5["object_size"], M5["object_shapes"], M5["object_max_r"])
    infos = line2dup.shape_infos(M5["angle_range"], [1.0], M5["angle_step"], 0.5)
    if n_angles is not None:
        infos = infos[:n_angles]
    return rotation_scale_templates(obj, mask, M5["class_id"], M5["num_features"], None, None, None, None, trainer=trainer, device=device,
                                    infos=infos)


def m5_frame(seed):
    rng = np.random.default_rng(seed)
    hw = M5["hw"]
    frame = dense_image(rng, hw, hw, 1, M5["frame_shapes"])[:, :, 0]
    obj, mask = textured_object(5, M5["object_size"], M5["object_shapes"], M5["object_max_r"])
    for _ in range(3):  # three instances at angles on the 0.5-degree grid
        centre = (int(rng.integers(300, hw - 300)), int(rng.integers(300, hw - 300)))
        paste_rotated(frame, obj, mask, centre, 0.5 * int(rng.integers(0, 720)), 1.0)
    return np.ascontiguousarray(frame)
