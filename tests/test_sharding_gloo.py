"""Template sharding over ranks, host logic on CPU: world_size-2 (and 3) gloo groups.

Each rank takes its slice of the golden candidate list of a fixture (the records its GPU would have produced
for its templates, with LOCAL template ids and in a scrambled order, as a device list is unordered), the ranks
exchange them with the same all_gather the GPU path uses, and sbm_merge_shard_records must give back the
reference's final list element by element.  No compute call is made: only the host side of the C ABI runs."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _local_records(case_name, rank, world, mode, n_templates):
    """the sbm_device_rec list rank `rank` would produce: its templates only, local ids, shuffled"""
    from _fixtures import load_case

    from shape_based_matching_b200 import DEVICE_REC
    from shape_based_matching_b200.sharding import shard_ids

    cand = load_case(case_name)["candidates"]
    ids = shard_ids(n_templates, rank, world, mode)
    local_of = {int(g): i for i, g in enumerate(ids)}
    rows = []
    for j, c in enumerate(cand):  # j is monotone within a template: a valid stand-in for the coarse scan index
        t = int(c["template_id"])
        if t in local_of:
            rows.append((int(c["x"]), int(c["y"]), float(c["similarity"]), local_of[t], j))
    recs = np.array(rows, dtype=DEVICE_REC) if rows else np.zeros(0, DEVICE_REC)
    np.random.default_rng(rank + 7).shuffle(recs)
    return recs, len(ids)


CASES = [("case2", 361), ("case1", 361), ("case0_1", 89)]


def _worker(rank, world, port, mode, cap, q):
    import torch.distributed as dist

    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        for case_name, n_templates in CASES:
            _one_case(rank, world, case_name, mode, n_templates, cap)
        q.put((rank, "ok"))
    except Exception as ex:  # surface the failure in the parent
        import traceback

        q.put((rank, traceback.format_exc() or repr(ex)))
    finally:
        dist.destroy_process_group()


def _one_case(rank, world, case_name, mode, n_templates, cap):
    import torch

    from _fixtures import load_case

    from shape_based_matching_b200 import DEVICE_REC
    from shape_based_matching_b200.sharding import allgather_records, merge_shard_records, shard_ids

    recs, _n_local = _local_records(case_name, rank, world, mode, n_templates)
    buf = np.zeros(cap * DEVICE_REC.itemsize, np.uint8)
    buf[: recs.nbytes] = recs.view(np.uint8)
    t_recs = torch.from_numpy(buf)
    t_cnt = torch.tensor([len(recs), 0], dtype=torch.int32)
    gathered, overflow = allgather_records(t_recs, t_cnt, cap)
    assert not overflow
    sizes = [[len(shard_ids(n_templates, r, world, mode))] for r in range(world)]
    gold = load_case(case_name)
    for finalize, want in ((False, gold["candidates"]), (True, gold["matches"])):
        got = merge_shard_records(gathered, sizes, mode, finalize)
        assert len(got) == len(want), (finalize, len(got), len(want))
        for k in ("x", "y", "template_id", "class_idx"):
            assert np.array_equal(got[k], want[k]), (finalize, k)
        assert np.array_equal(got["similarity"].view(np.uint32), want["similarity"].view(np.uint32))


@pytest.mark.parametrize("world,mode", [(2, 0), (2, 1), (3, 0)])
def test_template_shards_merge_to_the_reference_list(world, mode):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, 2048, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def test_shard_templates_partition_is_exact():
    from _fixtures import load_templates

    from shape_based_matching_b200.sharding import shard_ids, shard_templates

    _cid, tdesc, feats, _nf = load_templates("case2")
    for world in (1, 2, 4, 8):
        for mode in (0, 1):
            seen = np.zeros(tdesc.shape[0], int)
            total = 0
            for r in range(world):
                ids = shard_ids(tdesc.shape[0], r, world, mode)
                seen[ids] += 1
                td, ft = shard_templates(tdesc, feats, r, world, mode)
                assert np.array_equal(td, tdesc[ids])
                assert ft.shape[0] == int(td[:, :, 4].sum())
                total += ft.shape[0]
                # the first template's features are the global template's features
                if len(ids):
                    g0 = int(tdesc[: ids[0], :, 4].sum())
                    n0 = int(tdesc[ids[0], :, 4].sum())
                    assert np.array_equal(ft[:n0], feats[g0 : g0 + n0])
            assert (seen == 1).all() and total == feats.shape[0]


def test_merge_rejects_bad_input():
    from shape_based_matching_b200 import DEVICE_REC, SbmError
    from shape_based_matching_b200.sharding import merge_shard_records

    bad = np.zeros(1, DEVICE_REC)
    bad["template_gid"] = 5  # rank 0 holds only 2 templates
    with pytest.raises(SbmError):
        merge_shard_records([bad, np.zeros(0, DEVICE_REC)], [[2], [2]])
    assert len(merge_shard_records([np.zeros(0, DEVICE_REC)] * 2, [[2], [2]])) == 0


@pytest.mark.parametrize("world,mode", [(2, 0), (3, 1)])
def test_two_class_merge_follows_the_request_order(world, mode):
    """class_sizes covers EVERY installed class slot in slot order; the caller's class list (order, subsets) goes in
    separately, like Detector::match appends class by class in request order (line2Dup.cpp:1124-1140).  Expected lists:
    the reference's own two-class outputs (tests/golden/masked_two_class.npz)."""
    from _fixtures import load_masked_two_class

    from shape_based_matching_b200 import DEVICE_REC
    from shape_based_matching_b200.sharding import merge_shard_records, shard_ids

    z = load_masked_two_class()
    slots = sorted(c[0] for c in z["classes"])  # a_ring, circle
    n_of = {c[0]: c[1].shape[0] for c in z["classes"]}
    sizes = [[len(shard_ids(n_of[s], r, world, mode)) for s in slots] for r in range(world)]
    for tag, req in z["requests"]:
        order = [slots.index(c) for c in req] if req else None
        # the device lists hold each installed-and-matched class ONCE, whatever the request repeats
        uniq = sorted(set(order)) if order is not None else list(range(len(slots)))
        base = z["candidates_" + tag]
        per_rank = []
        for r in range(world):
            rows = []
            for k in uniq:
                ids = shard_ids(n_of[slots[k]], r, world, mode)
                local_of = {int(g): i for i, g in enumerate(ids)}
                gid0 = sum(sizes[r][:k])
                seen = set()
                for j, c in enumerate(base):
                    if int(c["class_idx"]) == k and int(c["template_id"]) in local_of:
                        key = (int(c["x"]), int(c["y"]), int(c["template_id"]), j)
                        rows.append((int(c["x"]), int(c["y"]), float(c["similarity"]), gid0 + local_of[int(c["template_id"])], j))
                        seen.add(key)
            recs = np.array(rows, dtype=DEVICE_REC) if rows else np.zeros(0, DEVICE_REC)
            np.random.default_rng(r).shuffle(recs)
            per_rank.append(recs)
        for finalize, want in ((False, base), (True, z["matches_" + tag])):
            got = merge_shard_records(per_rank, sizes, mode, finalize, class_order=order)
            assert len(got) == len(want), (tag, finalize, len(got), len(want))
            for f in ("x", "y", "template_id", "class_idx"):
                assert np.array_equal(got[f], want[f]), (tag, finalize, f)
            assert np.array_equal(got["similarity"].view(np.uint32), want["similarity"].view(np.uint32))
