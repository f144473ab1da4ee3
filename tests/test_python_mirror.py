"""Host logic of the Python mirror that needs no GPU: how Detector._batch hands a batch of frames to the C ABI (one
[N, H, W, C] block = an arithmetic sequence of addresses, a list = one address per frame) and how the packed result lists come
back (Detector._split).  The C entry point is replaced by a stand-in that records what it was given."""
import ctypes as C

import numpy as np

from shape_based_matching_b200 import detector as D
from shape_based_matching_b200._capi import MATCH_REC


class _NoLibrary(D.Detector):
    def __init__(self):  # no context: only the marshalling is exercised
        self._h = None

    def _class_args(self, class_ids):
        return None, 0, None


def _stand_in(seen, per_list):
    def fn(h, ptrs, mptrs, n, w, hh, c, stride, thr, arr, ncls, out, cap, begin):
        seen["ptrs"] = [ptrs[i] for i in range(n)]
        seen["shape"] = (n, hh, w, c)
        seen["masks"] = None if mptrs is None else [mptrs[i] for i in range(n)]
        recs = np.frombuffer((C.c_char * (cap * MATCH_REC.itemsize)).from_address(out.value), dtype=MATCH_REC)
        b = np.frombuffer(begin, dtype=np.int32, count=n + 1)
        b[:] = np.concatenate([[0], np.cumsum(per_list)])
        recs["x"][: b[n]] = np.arange(b[n])
        recs["template_id"][: b[n]] = np.repeat(np.arange(n), per_list)
        return 0
    return fn


def _check_lists(lists, per_list):
    assert [len(l) for l in lists] == list(per_list)
    at = 0
    for i, l in enumerate(lists):
        assert l.dtype == MATCH_REC
        assert np.array_equal(l["x"], np.arange(at, at + per_list[i])) and np.all(l["template_id"] == i)
        at += per_list[i]


def test_a_block_of_frames_is_an_arithmetic_sequence_of_addresses():
    det, seen = _NoLibrary(), {}
    frames = np.zeros((5, 24, 32, 3), np.uint8)
    per_list = [3, 0, 7, 1, 2]
    lists = det._batch(_stand_in(seen, per_list), frames, 90.0, [], None, None, 5)
    assert seen["shape"] == (5, 24, 32, 3) and seen["masks"] is None
    assert seen["ptrs"] == [frames.ctypes.data + i * 24 * 32 * 3 for i in range(5)]
    _check_lists(lists, per_list)
    gray = np.zeros((4, 16, 48), np.uint8)  # [N, H, W]: one-channel frames
    det._batch(_stand_in(seen, [0, 0, 0, 0]), gray, 90.0, [], None, None, 4)
    assert seen["shape"] == (4, 16, 48, 1) and seen["ptrs"] == [gray.ctypes.data + i * 16 * 48 for i in range(4)]


def test_a_list_of_frames_and_masks():
    det, seen = _NoLibrary(), {}
    frames = [np.full((24, 32, 3), i, np.uint8) for i in range(3)]
    frames[1] = np.asfortranarray(frames[1])  # not C-contiguous: copied, the copy's address is passed
    masks = [None, np.ones((24, 32), np.uint8), None]
    per_list = [2, 2, 0]
    lists = det._batch(_stand_in(seen, per_list), frames, 90.0, [], masks, None, 3)
    assert seen["shape"] == (3, 24, 32, 3)
    assert seen["ptrs"][0] == frames[0].ctypes.data and seen["ptrs"][2] == frames[2].ctypes.data
    assert seen["ptrs"][1] != frames[1].ctypes.data
    assert seen["masks"][0] is None and seen["masks"][1] is not None and seen["masks"][2] is None
    _check_lists(lists, per_list)
    assert det._batch(_stand_in(seen, []), [], 90.0, [], None, None, 0) == []
