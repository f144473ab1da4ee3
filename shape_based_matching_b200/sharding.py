"""Template sharding across the GPUs of one box (one process per GPU over torch.distributed).

Templates are independent given a frame's linear memories (the reference's `#pragma omp parallel for` over
templates, /root/reference/line2Dup.cpp:1166-1170), so rank r installs a slice of every class, every rank builds the
frame's linear memories locally, and the only exchange is an allgather of the compact record lists
(``sbm_device_rec``, 20 bytes each), merged by ``sbm_merge_shard_records`` into exactly the list the unsharded
``Detector::match`` returns.  Host logic only; works with the nccl backend (device tensors) and with gloo (CPU).
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import DEVICE_REC, MATCH_REC, check

INTERLEAVED, CONTIGUOUS = 0, 1


def shard_bounds(n, rank, world):
    """contiguous mode: [begin, end) of rank's block of n templates (blocks differ by at most one template)"""
    base, extra = divmod(n, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_ids(n, rank, world, mode=INTERLEAVED):
    """global template ids of a class of n templates owned by `rank` (ascending)"""
    if mode == INTERLEAVED:
        return np.arange(rank, n, world)
    b, e = shard_bounds(n, rank, world)
    return np.arange(b, e)


def shard_templates(tdesc, feats, rank, world, mode=INTERLEAVED):
    """slice of a class: tdesc[n, levels, 5] (last column = feature count), feats[sum nf, 3] ordered [template][level]
    -> (tdesc_local, feats_local) of the templates shard_ids() names, same ordering"""
    tdesc = np.asarray(tdesc)
    feats = np.asarray(feats)
    ids = shard_ids(tdesc.shape[0], rank, world, mode)
    per_t = tdesc[:, :, 4].sum(axis=1).astype(np.int64)
    begin = np.concatenate([[0], np.cumsum(per_t)])
    parts = [feats[begin[t] : begin[t + 1]] for t in ids]
    f = np.concatenate(parts) if parts else feats[:0]
    return np.ascontiguousarray(tdesc[ids]), np.ascontiguousarray(f)


def merge_shard_records(recs_per_rank, class_sizes, mode=INTERLEAVED, finalize=True, class_order=None):
    """recs_per_rank: list (rank order) of DEVICE_REC arrays; class_sizes[rank][class slot] = templates installed, for
    EVERY installed class slot in slot (= sorted id) order; class_order: the caller's class list as slot indices
    (repeats allowed; None = every slot once).  -> MATCH_REC array with GLOBAL template ids: the list the unsharded
    Detector::match returns."""
    L = _capi.lib()
    world = len(recs_per_rank)
    counts = np.array([len(r) for r in recs_per_rank], np.int32)
    allr = np.concatenate([np.ascontiguousarray(r, dtype=DEVICE_REC) for r in recs_per_rank]) if world else np.zeros(0, DEVICE_REC)
    allr = np.ascontiguousarray(allr)
    cs = np.ascontiguousarray(np.asarray(class_sizes, np.int32).reshape(world, -1))
    order = None if class_order is None else np.ascontiguousarray(class_order, np.int32)
    cap = max(1, int(counts.sum()) * (1 if order is None else max(1, len(order))))
    out = np.zeros(cap, MATCH_REC)
    n_out = C.c_int()
    check(L.sbm_merge_shard_records(allr.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p), world,
                                    cs.ctypes.data_as(C.c_void_p), cs.shape[1], int(mode),
                                    None if order is None else order.ctypes.data_as(C.c_void_p), 0 if order is None else len(order),
                                    int(bool(finalize)), out.ctypes.data_as(C.c_void_p), cap, C.byref(n_out)))
    return out[: n_out.value].copy()


def allgather_records(recs, count, cap, group=None):
    """The path's one exchange.  recs: uint8 tensor [cap * 20] (device tensor under nccl, CPU under gloo) holding
    this rank's sbm_device_rec list, count: int32 tensor [2] ([0] = records produced).  Returns (list of
    DEVICE_REC arrays in rank order, overflow flag) on the host of every rank."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    recs, count = recs.contiguous().view(-1), count.contiguous().view(-1)
    g_recs = torch.empty(world * recs.numel(), dtype=recs.dtype, device=recs.device)  # flat: gloo insists on it
    g_cnt = torch.empty(world * count.numel(), dtype=count.dtype, device=count.device)
    dist.all_gather_into_tensor(g_cnt, count, group=group)
    dist.all_gather_into_tensor(g_recs, recs, group=group)
    h_cnt = g_cnt.cpu().numpy().reshape(world, -1)
    h = g_recs.cpu().numpy().reshape(world, -1)
    out, overflow = [], False
    for r in range(world):
        n = int(h_cnt[r, 0])
        overflow |= n > cap
        out.append(h[r, : min(n, cap) * DEVICE_REC.itemsize].view(DEVICE_REC).copy())
    return out, overflow


def broadcast_unique_id(group=None):
    """rank 0's NCCL unique id (sbm_group_unique_id) on every rank.  torch.distributed is only the messenger here: the
    exchange of the match path itself runs inside the C ABI."""
    import torch.distributed as dist

    from .detector import Detector

    box = [Detector.group_unique_id() if dist.get_rank(group) == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return box[0]


class ShardedDetector:
    """Detector::match with the template set sharded over the ranks of a group (one process per GPU).  Same template
    store calls as Detector -- every rank is handed the FULL classes, the context keeps its slice -- and match()
    returns the full final list on every rank.  The allgather and the merge run inside the C ABI (sbm_group_match);
    torch.distributed only carries the 128-byte NCCL id at construction."""

    def __init__(self, num_features=63, T=(4, 8), weak_thresh=30.0, strong_thresh=60.0, device=0, group=None, mode=INTERLEAVED,
                 unique_id=None, rank=None, world=None):
        from .detector import Detector

        if unique_id is None:
            import torch.distributed as dist

            rank, world = dist.get_rank(group), dist.get_world_size(group)
            unique_id = broadcast_unique_id(group)
        self.rank, self.world, self.mode = int(rank), int(world), mode
        self.det = Detector(num_features, T, weak_thresh, strong_thresh, device=device)
        self.det.group_create(self.world, self.rank, unique_id, mode)
        self.device = device

    def setClassTemplates(self, class_id, tdesc, feats):
        self.det.setClassTemplates(class_id, tdesc, feats)

    def match(self, frame, threshold, class_ids=(), masks=None):
        """frame: HxW[xC] uint8 numpy array, the same on every rank -> global final MATCH_REC list"""
        return self.det.group_match_records(frame, threshold, class_ids, masks)

    def match_batch(self, frames, threshold, class_ids=()):
        return self.det.group_match_batch(frames, threshold, class_ids)

    def close(self):
        self.det.close()
