"""sbm_quantized_level1_static: the allocation-free entry point (pyramid level 1's quantised orientations with the
pyrDown'ed image in a static device array inside the library).  No counterpart in the reference; checked, through the C
ABI, against the existing entry point sbm_quantized_orientations(level = 1) -- bit for bit, the bar of every integer
stage here -- on two calls in a row with different images (the array keeps the previous call's contents).

Collected last on purpose (the file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _expected(det, img):
    _mag, _ang, q = det.quantized_orientations(img, level=1)
    return q


def _static(lib, d_img, w, h, c, weak, d_out, stream):
    from shape_based_matching_b200 import _capi

    _capi.check(lib.sbm_quantized_level1_static(C.c_void_p(d_img.data_ptr()), w, h, c, C.c_float(weak), C.c_void_p(d_out.data_ptr()),
                                                C.c_void_p(stream) if stream else None))


@pytest.mark.parametrize("channels", [3, 1])
def test_static_scratch_equals_the_allocating_entry_point(channels):
    import torch

    from _fixtures import load_case, to_gray
    from shape_based_matching_b200 import Detector, _capi

    lib = _capi.lib()
    case = load_case("case1")
    a = case["img"] if channels == 3 else to_gray(case["img"])[:, :, None]
    a = np.ascontiguousarray(a)
    rng = np.random.default_rng(7)
    # second image: another shape (smaller level-1 image than the first: stale bytes stay behind it in the array), odd size
    b = np.ascontiguousarray(np.roll(a, (37, 91), axis=(0, 1))[: a.shape[0] - 131, : a.shape[1] - 77])
    b = np.ascontiguousarray(np.clip(b.astype(np.int16) + rng.integers(-6, 7, b.shape), 0, 255).astype(np.uint8))
    det = Detector(63, (4, 8), 30.0, 60.0)
    st = torch.cuda.Stream()
    for img in (a, b, a):  # calls in a row, the third back on the first image
        h, w, c = img.shape
        want = _expected(det, img)
        d_img = torch.from_numpy(img).cuda()
        d_out = torch.full((h // 2, w // 2), 0xAA, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        _static(lib, d_img, w, h, c, 30.0, d_out, st.cuda_stream)
        st.synchronize()
        got = d_out.cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(got, want), "level-1 labels differ in %d pixels" % int((got != want).sum())
        assert (want != 0).sum() > 100  # the comparison is not between two empty images
    # NULL stream = the legacy default stream
    h, w, c = a.shape
    d_img = torch.from_numpy(a).cuda()
    d_out = torch.zeros((h // 2, w // 2), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    _static(lib, d_img, w, h, c, 30.0, d_out, 0)
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy(), _expected(det, a))
    det.close()


def test_static_scratch_limit_and_arguments():
    import torch

    from shape_based_matching_b200 import SbmError, _capi

    lib = _capi.lib()
    d = torch.zeros(16, dtype=torch.uint8, device="cuda")  # never touched: the calls below fail before any launch
    with pytest.raises(SbmError) as e:  # 4096 x 4096 x 3 -> 12 MiB of level-1 image, the array holds 3 MiB
        _static(lib, d, 4096, 4096, 3, 30.0, d, 0)
    assert e.value.code == -3 and "3145728" in str(e.value) and "12582912" in str(e.value)
    with pytest.raises(SbmError) as e:
        _static(lib, d, 64, 64, 2, 30.0, d, 0)
    assert e.value.code == -1
    with pytest.raises(SbmError) as e:
        _static(lib, d, 1, 64, 1, 30.0, d, 0)
    assert e.value.code == -1
    # the largest source the array holds is accepted: 2048 x 2048 x 3
    big = torch.zeros((2048, 2048, 3), dtype=torch.uint8, device="cuda")
    out = torch.ones((1024, 1024), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    _static(lib, big, 2048, 2048, 3, 30.0, out, 0)
    torch.cuda.synchronize()
    assert int(out.count_nonzero()) == 0  # a flat image has no gradient: every label 0
