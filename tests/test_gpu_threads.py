"""Threading contract of the C ABI ("Threading"): one context per thread runs concurrently, calls on
one shared context serialise on its mutex; either way every call returns the list a lone call returns."""
import threading

import numpy as np
import pytest

from _fixtures import load_case, load_templates

pytestmark = pytest.mark.gpu


def frames_of(case, n):
    base = case["img"]
    return [np.ascontiguousarray(np.roll(base, (7 * k, 11 * k), axis=(0, 1))) for k in range(n)]


def same(a, b):
    return len(a) == len(b) and all(np.array_equal(a[k], b[k]) for k in ("x", "y", "template_id", "class_idx")) and np.array_equal(
        a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))


def test_contexts_in_parallel_threads_and_one_shared_context():
    from shape_based_matching_b200 import Detector

    case = load_case("case0_1")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    frames = frames_of(case, 6)
    ref_det = Detector(nf, (4, 8))
    ref_det.setClassTemplates(cid, tdesc, feats)
    want = [ref_det.match_records(f, 85.0, [cid]) for f in frames]
    assert any(len(w) for w in want)

    errors = []

    def worker(det, order, rounds):
        try:
            for _ in range(rounds):
                for i in order:
                    got = det.match_records(frames[i], 85.0, [cid])
                    if not same(got, want[i]):
                        errors.append("frame %d differs (%d vs %d records)" % (i, len(got), len(want[i])))
                batch = det.match_batch([frames[i] for i in order], 85.0, [cid])
                for i, got in zip(order, batch):
                    if not same(got, want[i]):
                        errors.append("batch frame %d differs" % i)
        except Exception as e:  # noqa: BLE001 - reported below
            errors.append(repr(e))

    # one context per thread
    dets = []
    for _ in range(3):
        d = Detector(nf, (4, 8))
        d.setClassTemplates(cid, tdesc, feats)
        dets.append(d)
    threads = [threading.Thread(target=worker, args=(dets[k], list(np.roll(np.arange(6), k)), 3)) for k in range(3)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    # one context shared by two threads
    threads = [threading.Thread(target=worker, args=(dets[0], order, 2)) for order in ([0, 1, 2, 3, 4, 5], [5, 3, 1, 4, 2, 0])]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    for d in dets:
        d.close()
    ref_det.close()
