"""Edge cases of Detector::match on the GPU: the reference's asserts and limits (line2Dup.cpp:639, :751-752, :811,
:1195), unknown and multiple classes (:1124-1138), empty inputs, masks that remove everything, templates larger
than the frame (template_positions <= 0, :825), capacity growth, and shape changes on a live context.  Expectations
come from tests/_refcheck.py: the reference's own code (oracle/_ref) wherever the reference accepts the input (masks,
several classes, 8191 features, capacity growth, shape changes), the C++ restatement otherwise (frames smaller than
the refinement window and templates larger than the frame index out of bounds in the reference)."""
import numpy as np
import pytest

from _refcheck import Expected
from oracle import ref
from test_gpu_vs_port import synth_image, synth_templates

pytestmark = pytest.mark.gpu


def same_list(a, b):
    assert len(a) == len(b), (len(a), len(b))
    for k in ("x", "y", "template_id", "class_idx"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))


def make(rng, h=128, w=160, c=3, T=(4, 8), n=8, box=(40, 36), nf=24):
    from shape_based_matching_b200 import Detector

    img = synth_image(rng, h, w, c)
    tdesc, feats = synth_templates(rng, n, len(T), box, nf)
    det = Detector(nf, T, 30.0, 60.0)
    det.setClassTemplates("a", tdesc, feats)
    orc = Expected(T, 30.0, [("a", tdesc, feats)], num_features=nf)
    return det, orc, img, tdesc, feats


def test_frame_shapes_the_reference_asserts_on():
    from shape_based_matching_b200 import SbmError

    rng = np.random.default_rng(1)
    det, _orc, _img, _td, _ft = make(rng)
    for shape in ((100, 100, 3), (128, 164, 3), (72, 128, 1)):  # level-1 rows/cols not a multiple of T = 8
        with pytest.raises(SbmError) as e:
            det.match_records(np.zeros(shape, np.uint8), 50.0, ["a"])
        assert e.value.code == -1
    with pytest.raises(SbmError):
        det.match_records(np.zeros((128, 160, 2), np.uint8), 50.0, ["a"])  # 1 or 3 channels only
    with pytest.raises(SbmError):
        det.match_records(np.zeros((128, 160, 3), np.uint8), 50.0, ["a"], masks=np.zeros((64, 160), np.uint8))
    # the context is still usable afterwards
    assert isinstance(det.match_records(np.zeros((128, 160, 3), np.uint8), 50.0, ["a"]), np.ndarray)


def test_feature_count_limit_8191():
    from shape_based_matching_b200 import Detector, SbmError

    rng = np.random.default_rng(2)
    img = synth_image(rng, 256, 256, 1)
    for nf, ok in ((8191, True), (8192, False)):
        tdesc = np.array([[[200, 200, 0, 0, nf]]], np.int32)
        feats = np.stack([rng.integers(0, 200, nf), rng.integers(0, 200, nf), rng.integers(0, 8, nf)], 1).astype(np.int32)
        det = Detector(nf, (4,), 30.0, 60.0)
        det.setClassTemplates("big", tdesc, feats)
        if ok:
            orc = Expected((4,), 30.0, [("big", tdesc, feats)], num_features=nf).process(img)
            assert orc.pinned or not ref.available()
            same_list(det.match_records(img, 20.0, ["big"]), orc.match(20.0))
            assert np.array_equal(det.similarity_map("big", 0, (64, 64)), orc.similarity("big", 0))
        else:
            with pytest.raises(SbmError) as e:  # CV_Error "feature size too large" (line2Dup.cpp:1195)
                det.match_records(img, 20.0, ["big"])
            assert e.value.code == -1


def test_unknown_empty_and_multiple_classes():
    from shape_based_matching_b200 import Detector

    rng = np.random.default_rng(3)
    T = (4, 8)
    img = synth_image(rng, 128, 160, 3)
    ta, fa = synth_templates(rng, 6, 2, (40, 36), 24)
    tb, fb = synth_templates(rng, 5, 2, (30, 44), 30)
    det = Detector(30, T, 30.0, 60.0)
    assert len(det.match_records(img, 30.0)) == 0  # no templates at all
    det.setClassTemplates("b_second", tb, fb)
    det.setClassTemplates("a_first", ta, fa)
    orc = Expected(T, 30.0, [("a_first", ta, fa), ("b_second", tb, fb)], num_features=30).process(img)  # slots follow std::map order
    assert orc.pinned or not ref.available()
    both = det.match_records(img, 30.0)  # empty list = all classes in map order (line2Dup.cpp:1124-1130)
    same_list(both, orc.match(30.0))
    assert set(both["class_idx"]) == {0, 1}
    assert len(det.match_records(img, 30.0, ["nope"])) == 0  # unknown ids are skipped (:1136-1138)
    only_b = det.match_records(img, 30.0, ["nope", "b_second"])
    want_b = orc.match(30.0, ["nope", "b_second"])
    assert len(only_b) == len(want_b) and np.array_equal(only_b["x"], want_b["x"]) and np.array_equal(only_b["template_id"], want_b["template_id"])
    assert [m.class_id for m in det.match(img, 30.0, ["b_second"])][:1] in ([], ["b_second"])
    # caller's order matters for the candidate order: b before a
    same_list(det.match_candidates(img, 30.0, ["b_second", "a_first"]), orc.match(30.0, ["b_second", "a_first"], finalize=False))
    same_list(det.match_records(img, 30.0, ["b_second", "a_first"]), orc.match(30.0, ["b_second", "a_first"]))
    # a class named twice is matched twice (the reference appends per requested id, :1134-1139)
    same_list(det.match_candidates(img, 30.0, ["a_first", "a_first"]), orc.match(30.0, ["a_first", "a_first"], finalize=False))
    same_list(det.match_records(img, 30.0, ["a_first", "a_first"]), orc.match(30.0, ["a_first", "a_first"]))
    # removing a class
    det.setClassTemplates("a_first", ta[:0], fa[:0])
    assert set(det.match_records(img, 30.0)["class_idx"]) <= {0, 1}


def test_thresholds_and_capacity_growth():
    rng = np.random.default_rng(4)
    det, orc, img, _td, _ft = make(rng)
    orc.process(img)
    assert len(det.match_records(img, 100.5, ["a"])) == 0
    lots = orc.match(1.0)
    assert len(lots) > 50
    same_list(det.match_records(img, 1.0, ["a"], cap=4), lots)  # tiny output buffer: grown, not truncated
    same_list(det.match_candidates(img, 1.0, ["a"], cap=4), orc.match(1.0, finalize=False))
    same_list(det.match_records(img, 40.0, ["a"]), orc.match(40.0))  # and back to a selective threshold


def test_masks():
    rng = np.random.default_rng(5)
    det, orc, img, _td, _ft = make(rng)
    h, w = img.shape[:2]
    zero = np.zeros((h, w), np.uint8)
    orc.process(img, zero)
    assert orc.pinned or not ref.available()
    assert len(orc.match(5.0)) == 0
    assert len(det.match_records(img, 5.0, ["a"], masks=zero)) == 0
    assert not det.intermediate(0, 0).any()
    stripes = np.zeros((h, w), np.uint8)
    stripes[:, ::3] = 7  # any non-zero value keeps the pixel (copyTo(dst, mask), line2Dup.cpp:446-450)
    orc.process(img, stripes)
    same_list(det.match_records(img, 20.0, ["a"], masks=stripes), orc.match(20.0))
    full = np.full((h, w), 255, np.uint8)
    orc.process(img)
    same_list(det.match_records(img, 30.0, ["a"], masks=full), orc.match(30.0))


@pytest.mark.parametrize("hw", [(64, 96), (64, 128)], ids=["flat_level0", "tiled_level0"])
def test_templates_larger_than_the_frame_and_features_outside(hw):
    """level-0 W = 24 keeps the flat memories, W = 32 the tiled ones (k_refine_tiled): clamps that go negative, patches
    that run over the end of a row and features outside the image must read the same bytes in both"""
    from shape_based_matching_b200 import Detector

    rng = np.random.default_rng(6)
    T = (4, 8)
    img = synth_image(rng, hw[0], hw[1], 3)
    # template 0: bigger than the frame at both levels (template_positions <= 0); template 1: fits, but some of its
    # features lie outside the coarse image and are skipped (line2Dup.cpp:836); template 2: ordinary
    tdesc = np.array([[[400, 300, 0, 0, 20], [200, 150, 0, 0, 10]],
                      [[60, 40, 0, 0, 20], [30, 20, 0, 0, 10]],
                      [[30, 30, 0, 0, 20], [15, 15, 0, 0, 10]]], np.int32)
    feats = []
    for t in range(3):
        for l in range(2):
            bw, bh, nf = int(tdesc[t, l, 0]), int(tdesc[t, l, 1]), int(tdesc[t, l, 4])
            f = np.stack([rng.integers(0, bw + 1, nf), rng.integers(0, bh + 1, nf), rng.integers(0, 8, nf)], 1)
            if t == 1:
                f[:3, 0] += 500  # far outside the frame
            feats.append(f)
    feats = np.concatenate(feats).astype(np.int32)
    det = Detector(20, T, 30.0, 60.0)
    det.setClassTemplates("c", tdesc, feats)
    orc = Expected(T, 30.0, [("c", tdesc, feats)], num_features=20).process(img)
    assert not orc.pinned  # template 0 is larger than the frame: the reference's clamp indexes before its memories
    for thr in (10.0, 35.0):
        same_list(det.match_candidates(img, thr, ["c"]), orc.match(thr, finalize=False))
        same_list(det.match_records(img, thr, ["c"]), orc.match(thr))
    import os
    if os.environ.get("SBM_REFINE_TILED", "1") != "0":
        assert det.last_stats()["refine_tiled"] == (1 if hw[1] % 64 == 0 else 0)
    for t in range(3):
        assert np.array_equal(det.similarity_map("c", t, (hw[0] // 16, hw[1] // 16)), orc.similarity("c", t)), t


def test_shape_changes_on_a_live_context():
    rng = np.random.default_rng(7)
    det, orc, _img, _td, _ft = make(rng)
    for shape in ((128, 160, 3), (256, 192, 3), (64, 64, 1), (128, 160, 3), (256, 192, 1)):
        img = synth_image(rng, *shape)
        orc.process(img)
        same_list(det.match_records(img, 30.0, ["a"]), orc.match(30.0))
        frames = [img, np.ascontiguousarray(img[::-1]), img]
        got = det.match_batch(frames, 30.0, ["a"])
        for f, g in zip(frames, got):
            orc.process(f)
            same_list(g, orc.match(30.0))


# the last three: 2 to 4 rows left below the last full 16-row K1 tile (its vertical pass needs 5 rows beyond the tile)
@pytest.mark.parametrize("shape,T", [((32, 32, 1), (4, 8)), ((32, 48, 3), (4, 8)), ((16, 16, 3), (4,)), ((8, 64, 1), (2, 4)),
                                     ((36, 64, 1), (4,)), ((34, 80, 3), (2,)), ((52, 64, 3), (4,))])
def test_frames_smaller_than_a_kernel_tile(shape, T):
    """frames far smaller than any kernel tile: every intermediate, every raw similarity map and the (usually empty)
    lists still equal the restatement's"""
    from shape_based_matching_b200 import Detector

    rng = np.random.default_rng(sum(shape) + len(T))
    img = (rng.integers(0, 256, shape, dtype=np.uint8) // 4).astype(np.uint8)
    img[shape[0] // 4 : shape[0] // 2 + 2, shape[1] // 3 : shape[1] // 3 * 2] = 230  # one bright block: strong edges
    img[-5:, : shape[1] // 2] = 200  # and an edge in the last rows of the frame
    tdesc, feats = synth_templates(rng, 4, len(T), (10, 6), 8)
    det = Detector(8, T, 30.0, 60.0)
    det.setClassTemplates("s", tdesc, feats)
    det.set_debug(True)
    orc = Expected(T, 30.0, [("s", tdesc, feats)], num_features=8).process(img)
    for thr in (1.0, 30.0):
        same_list(det.match_candidates(img, thr, ["s"]), orc.match(thr, finalize=False))
        same_list(det.match_records(img, thr, ["s"]), orc.match(thr))
    for l in range(len(T)):
        for kind in range(4):
            assert np.array_equal(det.intermediate(l, kind), orc.intermediate(l, kind)), (l, kind)
    lv = len(T) - 1
    hw = ((shape[0] >> lv) // T[lv], (shape[1] >> lv) // T[lv])
    for t in range(4):
        assert np.array_equal(det.similarity_map("s", t, hw), orc.similarity("s", t)), t


def test_stage_overflow_on_the_device_path_raises_the_flag():
    """more than 65536 coarse candidates in one frame (threshold -1: every position of every template passes).  The host
    path grows its stage buffers and returns the complete list; the device-resident paths cannot synchronise, so they must
    raise the overflow flag (d_count[1]) instead of handing back a silently truncated, atomically ordered subset -- on the
    graph path, on the eager path (mask) and for batches -- until sbm_set_candidate_capacity makes room."""
    import torch

    from shape_based_matching_b200 import DEVICE_REC, Detector

    rng = np.random.default_rng(8)
    T = (4, 8)
    img = synth_image(rng, 512, 512, 1)
    tdesc, feats = synth_templates(rng, 80, 2, (40, 36), 12)
    det = Detector(12, T, 30.0, 60.0)
    det.setClassTemplates("a", tdesc, feats)
    orc = Expected(T, 30.0, [("a", tdesc, feats)], num_features=12).process(img)
    want = orc.match(-1.0, ["a"], finalize=False)
    assert len(want) > 65536
    h, w = img.shape[:2]
    d_img = torch.from_numpy(img).cuda()
    cap = 1 << 17
    recs = torch.zeros(cap * DEVICE_REC.itemsize, dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    full = np.full((h, w), 255, np.uint8)
    d_mask = torch.from_numpy(full).cuda()
    st = torch.cuda.Stream()

    def device_run(mask_ptr=0):
        cnt.zero_()
        torch.cuda.synchronize()
        det.match_device(d_img.data_ptr(), w, h, 1, -1.0, ["a"], d_records=recs.data_ptr(), cap=cap, d_count=cnt.data_ptr(),
                         stream=st.cuda_stream, d_mask=mask_ptr)
        st.synchronize()
        return [int(v) for v in cnt.cpu()]

    n, flag = device_run()
    assert flag == 1, "graph path: %d records reported without the overflow flag" % n
    n, flag = device_run(d_mask.data_ptr())
    assert flag == 1, "eager path: %d records reported without the overflow flag" % n
    # batches: every frame's flag
    frames = torch.stack([d_img, d_img])
    brecs = torch.zeros(2 * cap * DEVICE_REC.itemsize, dtype=torch.uint8, device="cuda")
    bcnt = torch.zeros((2, 2), dtype=torch.int32, device="cuda")
    det.match_batch_device(frames.data_ptr(), h * w, 2, w, h, 1, -1.0, ["a"], brecs.data_ptr(), cap, bcnt.data_ptr(), st.cuda_stream)
    st.synchronize()
    assert bcnt.cpu()[:, 1].tolist() == [1, 1]
    # the host path grows and returns everything, in the reference's order
    same_list(det.match_candidates(img, -1.0, ["a"], cap=1 << 17), want)
    # room made: the device path now returns the complete multiset with the flag down
    det.set_candidate_capacity(1 << 17)
    for mask_ptr in (0, d_mask.data_ptr()):
        n, flag = device_run(mask_ptr)
        assert flag == 0 and n == len(want)
        got = det.finalize_records(recs.cpu().numpy().view(DEVICE_REC)[:n], ["a"], finalize=False)
        same_list(got, want)
    # a caller buffer that is too small still raises it
    cnt.zero_()
    torch.cuda.synchronize()
    det.match_device(d_img.data_ptr(), w, h, 1, -1.0, ["a"], d_records=recs.data_ptr(), cap=1000, d_count=cnt.data_ptr(), stream=st.cuda_stream)
    st.synchronize()
    assert int(cnt.cpu()[1]) == 1 and int(cnt.cpu()[0]) == len(want)
