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


def merge_shard_records(recs_per_rank, class_sizes, mode=INTERLEAVED, finalize=True):
    """recs_per_rank: list (rank order) of DEVICE_REC arrays; class_sizes[rank][class slot] = templates installed.
    -> MATCH_REC array with GLOBAL template ids: the list the unsharded Detector::match returns."""
    L = _capi.lib()
    world = len(recs_per_rank)
    counts = np.array([len(r) for r in recs_per_rank], np.int32)
    allr = np.concatenate([np.ascontiguousarray(r, dtype=DEVICE_REC) for r in recs_per_rank]) if world else np.zeros(0, DEVICE_REC)
    allr = np.ascontiguousarray(allr)
    cs = np.ascontiguousarray(np.asarray(class_sizes, np.int32).reshape(world, -1))
    cap = max(1, int(counts.sum()))
    out = np.zeros(cap, MATCH_REC)
    n_out = C.c_int()
    check(L.sbm_merge_shard_records(allr.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p), world,
                                    cs.ctypes.data_as(C.c_void_p), cs.shape[1], int(mode), int(bool(finalize)),
                                    out.ctypes.data_as(C.c_void_p), cap, C.byref(n_out)))
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


class ShardedDetector:
    """Detector::match with the template set sharded over the ranks of a torch.distributed group.  Same template
    store calls as Detector; match() returns the full (global) final list on every rank."""

    def __init__(self, num_features=63, T=(4, 8), weak_thresh=30.0, strong_thresh=60.0, device=0, group=None, mode=INTERLEAVED,
                 rec_cap=8192):
        import torch.distributed as dist

        from .detector import Detector

        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.mode = mode
        self.rec_cap = rec_cap
        self.det = Detector(num_features, T, weak_thresh, strong_thresh, device=device)
        self.device = device
        self._sizes = {}  # class id -> global template count

    def setClassTemplates(self, class_id, tdesc, feats):
        tdesc = np.asarray(tdesc)
        self._sizes[class_id] = tdesc.shape[0]
        td, ft = shard_templates(tdesc, feats, self.rank, self.world, self.mode)
        self.det.setClassTemplates(class_id, td, ft)

    def class_sizes(self, class_ids):
        ids = list(class_ids) if class_ids else sorted(self._sizes)
        return [[len(shard_ids(self._sizes[c], r, self.world, self.mode)) for c in ids] for r in range(self.world)]

    def match_device(self, d_img, w, h, channels, threshold, class_ids, d_recs, d_cnt, stream=0):
        """enqueue this rank's shard on `stream`; results stay in the given device tensors"""
        self.det.match_device(d_img, w, h, channels, threshold, class_ids, d_records=d_recs.data_ptr(), cap=self.rec_cap,
                              d_count=d_cnt.data_ptr(), stream=stream)

    def match(self, frame, threshold, class_ids=()):
        """frame: HxWxC uint8 numpy array, the same on every rank -> global final MATCH_REC list"""
        import torch

        dev = torch.device("cuda", self.device)
        img = np.ascontiguousarray(frame)
        if img.ndim == 2:
            img = img[:, :, None]
        h, w, c = img.shape
        d_img = torch.from_numpy(img).to(dev)
        d_recs = torch.zeros(self.rec_cap * DEVICE_REC.itemsize, dtype=torch.uint8, device=dev)
        d_cnt = torch.zeros(2, dtype=torch.int32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        self.match_device(d_img.data_ptr(), w, h, c, threshold, class_ids, d_recs, d_cnt, stream)
        recs, overflow = allgather_records(d_recs, d_cnt, self.rec_cap, self.group)
        if overflow:
            raise _capi.SbmError(-3, "a shard produced more than rec_cap=%d records" % self.rec_cap)
        return merge_shard_records(recs, self.class_sizes(class_ids), self.mode, True)
