"""CPU: the C++ restatement (oracle/sbm_oracle.cpp) against the reference's OWN code (oracle/_ref) on the synthetic
configurations the GPU tests use -- masks, several classes, three pyramid levels, > 63 and 8191 features, ragged
widths.  Every configuration the reference accepts (tests/_refcheck.py: ref_admissible) must agree byte for byte at
every stage, so that the GPU tests, which compare against the same Expected object, are pinned to the reference and
not only to the builder's restatement."""
import numpy as np
import pytest

from _refcheck import Expected, ref_admissible, same_list
from oracle import ref
from test_gpu_vs_port import CONFIGS, synth_image, synth_templates

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")


def scenario(cfg):
    h, w, c, T, n, box, nf, thr, use_mask = cfg
    rng = np.random.default_rng(hash(cfg) & 0xFFFF)
    img = synth_image(rng, h, w, c)
    mask = None
    if use_mask:
        mask = np.full((h, w), 255, np.uint8)
        mask[: h // 3, : w // 2] = 0
    tdesc, feats = synth_templates(rng, n, len(T), box, nf)
    return img, mask, tdesc, feats


ADMISSIBLE = [i for i, cfg in enumerate(CONFIGS) if ref_admissible(cfg[:2], cfg[3], [("c", scenario(cfg)[2], None)])]


def test_enough_configurations_reach_the_reference():
    # masks (1, 15), three levels (3), > 63 features (2, 4), T = 8 on both levels (16), nibble / byte coarse layouts
    assert len(ADMISSIBLE) >= 9, ADMISSIBLE
    assert any(CONFIGS[i][8] for i in ADMISSIBLE), "no masked configuration is admissible"
    assert any(len(CONFIGS[i][3]) == 3 for i in ADMISSIBLE)


@pytest.mark.parametrize("i", ADMISSIBLE)
def test_restatement_equals_reference(i):
    cfg = CONFIGS[i]
    h, w, c, T, n, box, nf, thr, use_mask = cfg
    img, mask, tdesc, feats = scenario(cfg)
    exp = Expected(T, 30.0, [("c", tdesc, feats)], num_features=nf).process(img, mask)
    assert exp.pinned
    for l in range(len(T)):
        for kind in range(4):
            exp.intermediate(l, kind)  # asserts restatement == reference
    for t in range(n):
        exp.similarity("c", t)
    assert len(exp.match(thr, ["c"], finalize=False)) > 0
    exp.match(thr, ["c"], finalize=True)


def test_two_classes_and_request_order():
    rng = np.random.default_rng(3)
    T = (4, 8)
    img = synth_image(rng, 128, 160, 3)
    ta, fa = synth_templates(rng, 6, 2, (40, 36), 24)
    tb, fb = synth_templates(rng, 5, 2, (30, 44), 30)
    exp = Expected(T, 30.0, [("b_second", tb, fb), ("a_first", ta, fa)], num_features=30).process(img)
    assert exp.pinned
    both = exp.match(30.0)
    assert set(both["class_idx"]) == {0, 1}
    for req in (["b_second"], ["b_second", "a_first"], ["a_first", "b_second"], ["nope", "b_second"], ["a_first", "a_first"]):
        exp.match(30.0, req, finalize=False)
        exp.match(30.0, req, finalize=True)


def test_masks_against_the_reference():
    rng = np.random.default_rng(5)
    img = synth_image(rng, 128, 160, 3)
    tdesc, feats = synth_templates(rng, 8, 2, (40, 36), 24)
    h, w = img.shape[:2]
    stripes = np.zeros((h, w), np.uint8)
    stripes[:, ::3] = 7
    half = np.full((h, w), 255, np.uint8)
    half[:, : w // 2] = 0
    for mask, thr in ((np.zeros((h, w), np.uint8), 5.0), (stripes, 20.0), (half, 30.0), (np.full((h, w), 255, np.uint8), 30.0)):
        exp = Expected((4, 8), 30.0, [("a", tdesc, feats)], num_features=24).process(img, mask)
        assert exp.pinned
        for l in range(2):
            exp.intermediate(l, 0)
            exp.intermediate(l, 3)
        exp.match(thr, ["a"], finalize=False)
        exp.match(thr, ["a"])


def test_8191_features_against_the_reference():
    rng = np.random.default_rng(2)
    img = synth_image(rng, 256, 256, 1)
    nf = 8191
    tdesc = np.array([[[200, 200, 0, 0, nf]]], np.int32)
    feats = np.stack([rng.integers(0, 200, nf), rng.integers(0, 200, nf), rng.integers(0, 8, nf)], 1).astype(np.int32)
    exp = Expected((4,), 30.0, [("big", tdesc, feats)], num_features=nf).process(img)
    assert exp.pinned
    exp.similarity("big", 0)
    exp.match(20.0, ["big"])
    # two levels: the u16 path of both similarity() and similarityLocal()
    tdesc2 = np.array([[[200, 200, 0, 0, nf], [100, 100, 0, 0, 4095]]], np.int32)
    f1 = np.stack([rng.integers(0, 100, 4095), rng.integers(0, 100, 4095), rng.integers(0, 8, 4095)], 1).astype(np.int32)
    exp = Expected((4, 8), 30.0, [("big", tdesc2, np.concatenate([feats, f1]))], num_features=nf).process(synth_image(rng, 512, 512, 1))
    assert exp.pinned
    exp.similarity("big", 0)
    assert len(exp.match(6.0, ["big"], finalize=False)) > 0
    exp.match(6.0, ["big"])


def test_local_patches_against_the_reference():
    rng = np.random.default_rng(5)
    img = synth_image(rng, 256, 256, 3)
    tdesc, feats = synth_templates(rng, 3, 2, (60, 60), 90)
    exp = Expected((4, 8), 30.0, [("c", tdesc, feats)], num_features=90).process(img)
    assert exp.pinned
    for t in range(3):
        for (x, y) in ((32, 32), (100, 77), (163, 131)):
            exp.similarity_local("c", t, 0, x, y)
