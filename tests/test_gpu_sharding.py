"""Template sharding on the GPU: the G ranks of a box are emulated one after the other on cuda:0
(each with its own Detector holding its slice of the templates); the merged list must be the reference's golden
list for the unsharded template set.  Multi-rank plumbing (the all_gather) is covered on CPU by
tests/test_sharding_gloo.py and on real GPUs by `bench.py --gpus N`."""
import numpy as np
import pytest

from _fixtures import load_case, load_templates

pytestmark = pytest.mark.gpu


def same_list(a, b):
    assert len(a) == len(b), (len(a), len(b))
    for k in ("x", "y", "template_id", "class_idx"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32))


@pytest.mark.parametrize("name", ["case1", "case2", "case0_1"])
@pytest.mark.parametrize("world,mode", [(2, 0), (4, 1), (8, 0)])
def test_sharded_match_equals_reference(name, world, mode):
    import torch

    from shape_based_matching_b200 import DEVICE_REC, Detector
    from shape_based_matching_b200.sharding import merge_shard_records, shard_ids, shard_templates

    case = load_case(name)
    cid, tdesc, feats, nf = load_templates(case["templates"])
    thr = float(case["threshold"])
    img = np.ascontiguousarray(case["img"])
    h, w, c = img.shape
    d_img = torch.from_numpy(img).cuda()
    cap = 4096
    per_rank, sizes = [], []
    for r in range(world):
        td, ft = shard_templates(tdesc, feats, r, world, mode)
        det = Detector(nf, [int(t) for t in case["T"]], float(case["weak"]), float(case["strong"]))
        det.setClassTemplates(cid, td, ft)
        recs = torch.zeros(cap * DEVICE_REC.itemsize, dtype=torch.uint8, device="cuda")
        cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
        det.match_device(d_img.data_ptr(), w, h, c, thr, [cid], d_records=recs.data_ptr(), cap=cap, d_count=cnt.data_ptr(),
                         stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        n = int(cnt.cpu()[0])
        assert n <= cap
        per_rank.append(recs.cpu().numpy().view(DEVICE_REC)[:n].copy())
        sizes.append([len(shard_ids(tdesc.shape[0], r, world, mode))])
        det.close()
    same_list(merge_shard_records(per_rank, sizes, mode, finalize=False), case["candidates"])
    same_list(merge_shard_records(per_rank, sizes, mode, finalize=True), case["matches"])
