"""shape_based_matching::shapeInfo_producer of the drop-in header (include/line2Dup.h) against the reference's own class
(/root/reference/line2Dup.h:344-460 through oracle/_ref): the `infos` sequence element by element -- order (scale-major),
count and the float accumulation of `angle += angle_step` bit for bit -- and src_of / mask_of.  Host code: runs on CPU."""
import numpy as np
import pytest

from oracle import ref
from shape_based_matching_b200 import line2dup

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")

PARAMS = [
    ([0, 360], [1], 45.0, 0.5),            # test.cpp angle_test (test.cpp:281-286)
    ([0, 360], [1], 1.0, 0.5),             # the 1-degree set of test/case1
    ([], [0.8, 1.2], 15.0, 0.01),          # test.cpp scale_test: scales only
    ([0, 119.9], [0.9, 1.1], 0.1, 0.1),    # 0.1-degree steps x 3 scales = 3600 templates
    ([0, 360], [0.9, 1.1], 0.1, 0.1),      # float drift over 3600 steps: the last angle is 359.91275, not 359.9
    ([10], [0.5], 15.0, 0.5),              # one angle, one scale
    ([], [], 15.0, 0.5),                   # defaults: angle 0, scale 1
    ([-30, 30], [1, 2], 7.5, 0.25),
    ([0, 1], [1], 0.0002, 0.5),            # steps near eps * 10
]


@pytest.mark.parametrize("angle_range,scale_range,angle_step,scale_step", PARAMS)
def test_produce_infos_equals_reference(angle_range, scale_range, angle_step, scale_step):
    want = ref.shape_infos(angle_range, scale_range, angle_step, scale_step)
    got = line2dup.shape_infos(angle_range, scale_range, angle_step, scale_step)
    assert got.shape == want.shape, (got.shape, want.shape)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_m4_info_count():
    infos = line2dup.shape_infos([0, 119.9], [0.9, 1.1], 0.1, 0.1)
    assert infos.shape[0] == 3600 and len(np.unique(infos[:, 1])) == 3


@pytest.mark.parametrize("angle,scale", [(0.0, 1.0), (90.0, 1.0), (180.0, 0.9), (270.0, 1.1), (45.0, 0.8), (90.0005, 0.5), (0.0, 1.1)])
def test_src_of_and_mask_of_equal_reference(angle, scale):
    rng = np.random.default_rng(int(angle * 10 + scale * 100))
    img = rng.integers(0, 256, (75, 101, 3), dtype=np.uint8)
    gray = rng.integers(0, 256, (64, 48), dtype=np.uint8)
    mask = (rng.random((75, 101)) > 0.3).astype(np.uint8) * 255
    for src in (img, gray):
        assert np.array_equal(line2dup.shape_transform(src, angle, scale), ref.shape_transform(src, angle, scale))
    assert np.array_equal(line2dup.shape_transform(mask, angle, scale, as_mask=True), ref.shape_transform(mask, angle, scale, as_mask=True))
