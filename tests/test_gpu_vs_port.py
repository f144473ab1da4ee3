"""GPU parity on seeded synthetic inputs: shapes, channel counts, T sets, pyramid depths and masks the fixtures do
not cover.  Expectations come from tests/_refcheck.py: the reference's OWN code (oracle/_ref) for every configuration
the reference accepts (all but two of those below), with the C++ restatement (oracle/sbm_oracle.cpp) run beside it and
required to agree; the restatement alone only for the shapes the reference asserts on (odd coarse widths).
Bar: byte-exact intermediates, identical candidate and final lists."""
import numpy as np
import pytest

from _refcheck import Expected
from oracle import ref

pytestmark = pytest.mark.gpu


def synth_image(rng, h, w, c):
    import cv2

    img = cv2.GaussianBlur(rng.integers(0, 256, (h, w, c), dtype=np.uint8), (5, 5), 0).reshape(h, w, c) // 3
    for _ in range(6):  # bright polygons and ellipses: strong, coherent edges
        col = [int(v) for v in rng.integers(120, 256, c)]
        if rng.random() < 0.5:
            pts = rng.integers(0, [w, h], (4, 2)).astype(np.int32)
            cv2.fillPoly(img, [pts], col)
        else:
            cv2.ellipse(img, (int(rng.integers(0, w)), int(rng.integers(0, h))), (int(rng.integers(8, w // 3)), int(rng.integers(8, h // 3))),
                        float(rng.integers(0, 180)), 0, 360, col, -1)
    return np.ascontiguousarray(img)


def synth_templates(rng, n, levels, box, nf0):
    """n template pyramids: random features inside a box that halves per level (like cropTemplates' output)"""
    tdesc = np.zeros((n, levels, 5), np.int32)
    feats = []
    for t in range(n):
        for l in range(levels):
            bw, bh = max(box[0] >> l, 4), max(box[1] >> l, 4)
            nf = max(nf0 >> l, 3)
            tdesc[t, l] = (bw, bh, 0, 0, nf)
            f = np.stack([rng.integers(0, bw + 1, nf), rng.integers(0, bh + 1, nf), rng.integers(0, 8, nf)], 1)
            feats.append(f)
    return tdesc, np.concatenate(feats).astype(np.int32)


CONFIGS = [
    # h, w, c, T, n_templates, box, nf, threshold, use_mask
    (128, 160, 3, (4, 8), 12, (40, 36), 24, 35.0, False),
    (128, 160, 1, (4, 8), 12, (40, 36), 24, 35.0, True),
    (96, 96, 1, (4,), 6, (30, 30), 70, 30.0, False),       # single level, > 63 features (u16 accumulation path)
    (256, 192, 3, (4, 4, 8), 8, (64, 48), 40, 30.0, False),  # three pyramid levels
    (192, 320, 3, (8, 8), 5, (100, 60), 200, 30.0, False),   # wide template, many features
    (64, 64, 3, (2, 4), 4, (16, 16), 12, 30.0, False),       # T = 2 at level 0 (rows not a multiple of 4 bytes * 3)
    (64, 66, 3, (2,), 4, (20, 16), 16, 30.0, False),         # width not a multiple of 4: unaligned rows everywhere
    (48, 70, 1, (2,), 4, (20, 16), 16, 30.0, True),          # same, one channel, masked
    (192, 208, 3, (4, 8), 6, (60, 50), 50, 30.0, False),     # width not a multiple of the 64 / 128 pixel kernel tiles
    # the coarsest level's memories are nibble-packed when its W is even (k_similarity_nib), byte-wide otherwise:
    (64, 72, 1, (2,), 4, (20, 16), 16, 30.0, False),         # T = 2 on top, W = 36: packed, generic writer path
    (96, 128, 3, (16,), 4, (40, 36), 20, 30.0, False),       # T = 16 on top, W = 8: packed, generic writer path
    (80, 144, 1, (4, 8), 5, (36, 30), 30, 30.0, True),       # T = 8 on top, W = 9 (odd): byte-wide fallback
    (128, 176, 3, (4, 8), 5, (40, 36), 90, 30.0, False),     # W = 11 (odd) with > 63 features: byte-wide, u16 widening
    (128, 224, 3, (4, 8), 5, (40, 36), 90, 30.0, False),     # W = 14 (W % 8 = 6) with > 63 features: packed, ragged row ends
    (88, 104, 3, (4, 4), 6, (30, 26), 40, 25.0, False),      # level-0 W = 26 (W % 4 = 2): k_refine's per-row shift path
    (128, 192, 3, (4, 8), 6, (40, 36), 40, 30.0, True),      # level-0 W = 48: tiled level 0 (k_refine_tiled) with a mask
    (192, 256, 3, (8, 8), 5, (60, 50), 70, 10.0, False),     # T = 8 on both levels, level-0 W = 32: tiled writer for T = 8
    # the coarsest level is stored as hit / any bit planes when its W is a multiple of 32 (k_similarity_bits):
    (256, 512, 3, (4, 8), 8, (64, 48), 40, 30.0, False),     # T = 8 on top, W = 32 (16-cell ballots), < 128 features: 7-plane counters
    (64, 128, 1, (4,), 6, (30, 30), 70, 30.0, True),         # T = 4 on top, W = 32 (whole-word ballots), masked
    (48, 128, 3, (2,), 4, (20, 16), 16, 30.0, False),        # T = 2 on top, W = 64 (two words per cell row)
    (64, 512, 1, (16,), 4, (60, 30), 20, 25.0, False),       # T = 16 on top, W = 32 (8-cell ballots)
    (256, 512, 3, (4, 8), 4, (90, 70), 600, 20.0, False),    # 300 coarse features: 10-plane counters, odd / even parity classes
    (512, 512, 1, (4, 8), 3, (200, 200), 4000, 10.0, False), # 2000 coarse features: 13-plane counters, two offset chunks
    (128, 256, 3, (4,), 5, (40, 36), 24, 30.0, False),       # W = 64 = exactly one warp of 64-position threads per row pair
]


@pytest.mark.parametrize("cfg", CONFIGS, ids=[str(i) for i in range(len(CONFIGS))])
def test_cuda_equals_port(cfg):
    from shape_based_matching_b200 import Detector

    h, w, c, T, n, box, nf, thr, use_mask = cfg
    rng = np.random.default_rng(hash(cfg) & 0xFFFF)
    img = synth_image(rng, h, w, c)
    mask = None
    if use_mask:
        mask = np.full((h, w), 255, np.uint8)
        mask[: h // 3, : w // 2] = 0
    tdesc, feats = synth_templates(rng, n, len(T), box, nf)
    det = Detector(nf, T, 30.0, 60.0)
    det.setClassTemplates("c", tdesc, feats)
    det.set_debug(True)
    orc = Expected(T, 30.0, [("c", tdesc, feats)], num_features=nf).process(img, mask)
    if ref.available():  # which configurations reach the reference's own code is fixed (tests/test_port_vs_ref_synthetic.py)
        assert orc.pinned == (CONFIGS.index(cfg) not in (11, 14)), "configuration %d: reference admissibility changed" % CONFIGS.index(cfg)
    got = det.match_records(img, thr, ["c"], masks=mask)
    for l in range(len(T)):
        for kind in range(4):
            a, b = det.intermediate(l, kind), orc.intermediate(l, kind)
            bad = np.flatnonzero(a != b)
            assert bad.size == 0, "level %d kind %d: %d bytes differ, first at %s" % (l, kind, bad.size, bad[:5])
    hw = ((h >> (len(T) - 1)) // T[-1], (w >> (len(T) - 1)) // T[-1])
    for t in range(n):
        assert np.array_equal(det.similarity_map("c", t, hw), orc.similarity("c", t)), "raw similarity map of template %d" % t
    cand, want_cand = det.match_candidates(img, thr, ["c"], masks=mask), orc.match(thr, ["c"], finalize=False)
    assert len(want_cand) > 0, "config produces no candidates: not a test"
    for name, a, b in (("candidates", cand, want_cand), ("final", got, orc.match(thr, ["c"], finalize=True))):
        assert len(a) == len(b), "%s: %d vs %d" % (name, len(a), len(b))
        for k in ("x", "y", "template_id"):
            assert np.array_equal(a[k], b[k]), (name, k)
        assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32)), name


def test_local_patch_sums():
    from shape_based_matching_b200 import Detector

    rng = np.random.default_rng(5)
    img = synth_image(rng, 256, 256, 3)
    tdesc, feats = synth_templates(rng, 3, 2, (60, 60), 90)
    det = Detector(90, (4, 8))
    det.setClassTemplates("c", tdesc, feats)
    det.match_records(img, 50.0, ["c"])
    orc = Expected((4, 8), 30.0, [("c", tdesc, feats)], num_features=90).process(img)
    assert orc.pinned or not ref.available()
    for t in range(3):
        for (x, y) in ((32, 32), (100, 77), (163, 131)):
            assert np.array_equal(det.similarity_local("c", t, 0, x, y), orc.similarity_local("c", t, 0, x, y))


def test_flat_byte_wide_layouts_by_environment():
    """SBM_COARSE_BITS=0 keeps k_similarity_nib (two responses per byte) where the default is bit planes;
    SBM_COARSE_NIBBLES=0 on top keeps k_similarity (one byte per response) on the coarsest level and SBM_REFINE_TILED=0
    keeps k_refine (flat memories) on the refined levels for every shape: the same parity tests in fresh processes,
    including the > 1024-offset chunks of the 8191-feature template"""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # 1: no bit planes -> the nibble-packed layout wherever W is even (M1, configurations 17-23); 2: bytes and flat everywhere
    for extra in (dict(SBM_COARSE_BITS="0"), dict(SBM_COARSE_BITS="0", SBM_COARSE_NIBBLES="0", SBM_REFINE_TILED="0")):
        env = dict(os.environ, **extra)
        r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "tests/test_gpu_vs_port.py::test_cuda_equals_port",
                            "tests/test_gpu_edge_cases.py::test_feature_count_limit_8191",
                            "tests/test_gpu_edge_cases.py::test_templates_larger_than_the_frame_and_features_outside", "tests/test_gpu_parity.py"],
                           cwd=root, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
        assert r.returncode == 0, (extra, r.stdout[-3000:])


def test_orientation_kernels_by_environment():
    """K1 exists twice: k_orient_mma (blur and Sobel as int8 matrix products on the tensor cores, the default) and k_orient
    (integer pipe, SBM_ORIENT_MMA=0).  Both must produce the reference's quantised images byte for byte, k_orient_mma with its
    source window staged by the copy engine (cp.async.bulk.tensor, the default for 16-byte aligned rows) and by plain loads
    (SBM_ORIENT_TMA=0), with one tile per CTA and with two (both windows requested at once: the default only on large levels, SBM_ORIENT_TILES=1 / 2),
    and at every tile height it is instantiated for (SBM_ORIENT_TH: 12, 20, 28, 36 rows) -- the same parity tests in fresh processes:
    fixtures colour and gray (intermediates byte-exact), the synthetic configurations (odd and ragged widths, three levels,
    masks) and the edge cases."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for extra in (dict(SBM_ORIENT_MMA="0"), dict(SBM_ORIENT_TMA="0"), dict(SBM_ORIENT_TH="12"), dict(SBM_ORIENT_TH="20"), dict(SBM_ORIENT_TH="28"),
                  dict(SBM_ORIENT_TILES="1"), dict(SBM_ORIENT_TILES="2")):
        env = dict(os.environ, **extra)
        r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "tests/test_gpu_vs_port.py::test_cuda_equals_port",
                            "tests/test_gpu_edge_cases.py::test_masks", "tests/test_gpu_edge_cases.py::test_frames_smaller_than_a_kernel_tile",
                            "tests/test_gpu_edge_cases.py::test_frame_shapes_the_reference_asserts_on", "tests/test_gpu_parity.py"],
                           cwd=root, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
        assert r.returncode == 0, (extra, r.stdout[-3000:])
