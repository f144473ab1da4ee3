"""The multi-GPU group behind the C ABI (sbm_group_*: NCCL bound at run time, one allgather of result blocks).  On one
GPU a group of ONE rank runs the whole code path -- communicator, slice bookkeeping, block export, ncclAllGather,
merge / finalize -- and must reproduce the reference's lists; with >= 2 GPUs (gpurun --gpus 2) two processes shard case1's
templates / frames for real (tests/cpp/group_driver.cpp drives the C++ Detector::setGroup the same way)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from _fixtures import load_case, load_masked_two_class, load_templates

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def same_list(a, b):
    assert len(a) == len(b), (len(a), len(b))
    for k in ("x", "y", "template_id", "class_idx"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))


@pytest.mark.parametrize("mode", [0, 1, 2], ids=["templates", "templates_contiguous", "frames"])
def test_group_of_one_equals_reference(mode):
    from shape_based_matching_b200 import Detector

    case = load_case("case1")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    det = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]))
    det.group_create(1, 0, Detector.group_unique_id(), mode)
    det.setClassTemplates(cid, tdesc, feats)
    thr = float(case["threshold"])
    frames = [case["img"], np.ascontiguousarray(case["img"][:, ::-1]), case["img"]]
    lists = det.group_match_batch(frames, thr, [cid])
    assert len(lists) == 3
    same_list(lists[0], case["matches"])
    same_list(lists[2], case["matches"])
    if mode != 2:
        same_list(det.group_match_records(case["img"], thr, [cid]), case["matches"])
    # device-resident + finalize
    import torch

    d = torch.from_numpy(np.stack(frames)).cuda()
    cap = 256
    blk = Detector.group_block_bytes(cap)
    g = torch.zeros(3 * blk, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        det.group_match_batch_device(d.data_ptr(), d[0].numel(), 3, d.shape[2], d.shape[1], d.shape[3], thr, [cid], g.data_ptr(), cap, st.cuda_stream)
    st.synchronize()
    again = det.group_finalize(g.cpu().numpy(), 3, cap, [cid])
    for a, b in zip(again, lists):
        same_list(a, b)
    # a block too small for the frame raises, it does not truncate
    tiny = torch.zeros(3 * Detector.group_block_bytes(2), dtype=torch.uint8, device="cuda")
    det.group_match_batch_device(d.data_ptr(), d[0].numel(), 3, d.shape[2], d.shape[1], d.shape[3], thr, [cid], tiny.data_ptr(), 2, 0)
    torch.cuda.synchronize()
    from shape_based_matching_b200 import SbmError

    with pytest.raises(SbmError) as e:
        det.group_finalize(tiny.cpu().numpy(), 3, 2, [cid])
    assert e.value.code == -3
    det.close()


def test_group_masks_and_two_classes():
    from shape_based_matching_b200 import Detector

    z = load_masked_two_class()
    det = Detector(int(z["num_features"]), [int(t) for t in z["T"]], float(z["weak"]), float(z["strong"]))
    det.group_create(1, 0, Detector.group_unique_id(), 0)
    for cid, td, ft in z["classes"]:
        det.setClassTemplates(cid, td, ft)
    thr = float(z["threshold"])
    for tag, req in z["requests"]:
        same_list(det.group_match_records(z["img"], thr, req), z["matches_" + tag])
        same_list(det.group_match_records(z["img"], thr, req, masks=z["mask"]), z["masked_matches_" + tag])
    det.close()


def test_group_grows_its_blocks():
    """more records per frame than the group's initial block holds (256): every rank grows alike and re-runs"""
    from _refcheck import Expected
    from shape_based_matching_b200 import Detector
    from test_gpu_vs_port import synth_image, synth_templates

    rng = np.random.default_rng(11)
    img = synth_image(rng, 256, 256, 3)
    tdesc, feats = synth_templates(rng, 12, 2, (40, 36), 24)
    exp = Expected((4, 8), 30.0, [("a", tdesc, feats)], num_features=24).process(img)
    want = exp.match(1.0, ["a"])
    assert len(want) > 256
    det = Detector(24, (4, 8))
    det.group_create(1, 0, Detector.group_unique_id(), 0)
    det.setClassTemplates("a", tdesc, feats)
    same_list(det.group_match_records(img, 1.0, ["a"], cap=1 << 16), want)
    det.close()


@pytest.mark.parametrize("mode", [0, 2], ids=["templates", "frames"])
def test_group_batch_longer_than_a_chunk(mode):
    """the host-side group call feeds a batch in chunks of 32 frames (exchange + host ordering of chunk k overlap chunk
    k+1): 70 distinct frames = chunks of 32, 32 and 6 must come back in order, equal to one blocking match per frame"""
    from shape_based_matching_b200 import Detector

    case = load_case("case1")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    thr = float(case["threshold"])
    frames = [np.ascontiguousarray(np.roll(case["img"], (8 * i, 16 * i), axis=(0, 1))) for i in range(70)]
    single = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]))
    single.setClassTemplates(cid, tdesc, feats)
    want = [single.match_records(f, thr, [cid]) for f in frames]
    single.close()
    assert len({len(w) for w in want}) > 1  # the frames really differ
    det = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]))
    det.group_create(1, 0, Detector.group_unique_id(), mode)
    det.setClassTemplates(cid, tdesc, feats)
    for _ in range(2):  # the second pass reuses the chunk events, the pool and the pinned blocks
        lists = det.group_match_batch(frames, thr, [cid])
        assert len(lists) == 70
        for a, b in zip(lists, want):
            same_list(a, b)
    # an output buffer that is too small reports the size it needs (the mirror retries with it)
    lists = det.group_match_batch(frames, thr, [cid], cap=8)
    for a, b in zip(lists, want):
        same_list(a, b)
    det.close()


WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
from _fixtures import load_case, load_templates
from shape_based_matching_b200 import Detector
rank, world, mode = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
uid = open(sys.argv[4], "rb").read()
case = load_case("case1")
cid, tdesc, feats, nf = load_templates(case["templates"])
det = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]), device=rank)
det.group_create(world, rank, uid, mode)
det.setClassTemplates(cid, tdesc, feats)
thr = float(case["threshold"])
def same(a, b):
    assert len(a) == len(b), (len(a), len(b))
    for k in ("x", "y", "template_id", "class_idx"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))
flip = np.ascontiguousarray(case["img"][:, ::-1])
if mode != 2:
    assert det.numTemplates() == tdesc.shape[0]
    same(det.group_match_records(case["img"], thr, [cid]), case["matches"])
    lists = det.group_match_batch([case["img"], flip, case["img"]], thr, [cid])
    assert len(lists) == 3
    same(lists[0], case["matches"]); same(lists[2], case["matches"])
else:
    mine = [case["img"], flip] if rank %% 2 == 0 else [flip, case["img"]]
    lists = det.group_match_batch(mine, thr, [cid])
    assert len(lists) == 2 * world
    for r in range(world):
        same(lists[2 * r + (0 if r %% 2 == 0 else 1)], case["matches"])
# a batch longer than one chunk of the host-side group call (32 frames): 40 frames per rank, every frame different
single = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]), device=rank)
single.setClassTemplates(cid, tdesc, feats)
def frame(i):
    return np.ascontiguousarray(np.roll(case["img"], (8 * i, 16 * i), axis=(0, 1)))
if mode != 2:
    fr = [frame(i) for i in range(40)]
    lists = det.group_match_batch(fr, thr, [cid])
    assert len(lists) == 40
    for f, l in zip(fr, lists):
        same(l, single.match_records(f, thr, [cid]))
else:
    lists = det.group_match_batch([frame(40 * rank + i) for i in range(40)], thr, [cid])
    assert len(lists) == 40 * world
    for r in range(world):
        for i in range(0, 40, 3):
            same(lists[40 * r + i], single.match_records(frame(40 * r + i), thr, [cid]))
print("rank %%d ok" %% rank)
'''


@pytest.mark.parametrize("mode", [0, 1, 2], ids=["templates", "templates_contiguous", "frames"])
def test_two_processes_two_gpus(mode, tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    from shape_based_matching_b200 import Detector

    uid = tmp_path / "uid.bin"
    uid.write_bytes(Detector.group_unique_id())
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", str(mode), str(uid)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                              text=True) for r in range(2)]
    for r, p in enumerate(procs):
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0 and ("rank %d ok" % r) in out, out[-3000:]
