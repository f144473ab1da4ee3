"""Python mirror of ``line2Dup::Detector`` (/root/reference/line2Dup.h:257-333) over the C ABI.

Same method names and argument meaning as the reference class; every pipeline stage of ``match`` runs in
the sm_100a CUDA library (``libsbm_b200.so``).  This module holds host bookkeeping only: the template
store (``std::map<std::string, std::vector<TemplatePyramid>>`` -> sorted dict) and ctypes marshalling.
"""
import ctypes as C
from collections import namedtuple

import numpy as np

from . import _capi
from ._capi import DEVICE_REC, FEATURE, MATCH_REC, TEMPLATE_DESC, check

Match = namedtuple("Match", "x y similarity class_id template_id")  # line2Dup.h:222-255
Feature = namedtuple("Feature", "x y label")  # line2Dup.h:116-129
Template = namedtuple("Template", "width height tl_x tl_y pyramid_level features")  # line2Dup.h:131-153


class Detector:
    def __init__(self, num_features=63, T=(4, 8), weak_thresh=30.0, strong_thresh=60.0, device=0):
        # Detector::Detector(int, std::vector<int>, float, float)  line2Dup.cpp:1071-1076 (defaults :1056-1062)
        self._L = _capi.lib()
        self.T_at_level = [int(t) for t in T]
        self.pyramid_levels = len(self.T_at_level)
        self.num_features = int(num_features)
        self.weak_threshold = float(weak_thresh)
        self.strong_threshold = float(strong_thresh)
        self.device = int(device)
        self._classes = {}  # class_id -> (tdesc[n,levels] TEMPLATE_DESC, feats FEATURE)
        self._slots = []  # sorted class ids as uploaded
        self._dirty = True
        h = C.c_void_p()
        arr = (C.c_int * self.pyramid_levels)(*self.T_at_level)
        check(self._L.sbm_create(arr, self.pyramid_levels, self.weak_threshold, self.strong_threshold, self.num_features,
                                 self.device, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.sbm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- template store -------------------------------------------------------------------------
    def getT(self, pyramid_level):
        return self.T_at_level[pyramid_level]

    def pyramidLevels(self):
        return self.pyramid_levels

    def classIds(self):
        return sorted(self._classes)

    def numClasses(self):
        return len(self._classes)

    def numTemplates(self, class_id=None):
        if class_id is None:
            return sum(v[0].shape[0] for v in self._classes.values())
        return self._classes[class_id][0].shape[0] if class_id in self._classes else 0

    def setClassTemplates(self, class_id, tdesc, feats):
        """Install a whole class from flat arrays: tdesc[n, levels, 5] = (width, height, tl_x, tl_y, nf),
        feats[sum nf, 3] = (x, y, label), ordered [template][level] -- what readClass builds
        (line2Dup.cpp:1507-1546)."""
        tdesc = np.asarray(tdesc)
        if tdesc.dtype != TEMPLATE_DESC:
            t = np.zeros(tdesc.shape[:2], TEMPLATE_DESC)
            for i, k in enumerate(TEMPLATE_DESC.names):
                t[k] = tdesc[..., i]
            tdesc = t
        feats = np.asarray(feats)
        if feats.dtype != FEATURE:
            f = np.zeros(feats.shape[0], FEATURE)
            for i, k in enumerate(FEATURE.names):
                f[k] = feats[:, i]
            feats = f
        if tdesc.ndim != 2 or tdesc.shape[1] != self.pyramid_levels:
            raise ValueError("tdesc must be [n_templates, pyramid_levels]")
        if int(tdesc["num_features"].sum()) != feats.shape[0]:
            raise ValueError("feature count mismatch")
        self._classes[class_id] = (np.ascontiguousarray(tdesc), np.ascontiguousarray(feats))
        self._dirty = True

    def getTemplates(self, class_id, template_id):
        # CV_Assert on unknown class / id (line2Dup.cpp:1452-1458)
        if class_id not in self._classes:
            raise KeyError(class_id)
        tdesc, feats = self._classes[class_id]
        if not 0 <= template_id < tdesc.shape[0]:
            raise IndexError(template_id)
        start = int(tdesc["num_features"].ravel()[: template_id * self.pyramid_levels].sum())
        out = []
        for l in range(self.pyramid_levels):
            d = tdesc[template_id, l]
            nf = int(d["num_features"])
            f = feats[start : start + nf]
            out.append(Template(int(d["width"]), int(d["height"]), int(d["tl_x"]), int(d["tl_y"]), l,
                                [Feature(int(a["x"]), int(a["y"]), int(a["label"])) for a in f]))
            start += nf
        return out

    def _sync_templates(self):
        if not self._dirty:
            return
        check(self._L.sbm_clear_templates(self._h))
        self._slots = sorted(self._classes)  # std::map iteration order (line2Dup.cpp:1127-1129)
        for k, cid in enumerate(self._slots):
            tdesc, feats = self._classes[cid]
            check(self._L.sbm_set_class_templates(self._h, k, tdesc.shape[0], tdesc.ctypes.data_as(C.c_void_p),
                                                  feats.ctypes.data_as(C.c_void_p)))
        self._dirty = False

    def _class_args(self, class_ids):
        self._sync_templates()
        if not class_ids:
            return None, 0, None
        idx = [self._slots.index(c) if c in self._slots else -1 for c in class_ids]
        arr = (C.c_int * len(idx))(*idx)
        return arr, len(idx), arr

    # ---- match ----------------------------------------------------------------------------------
    @staticmethod
    def _img(source):
        img = np.asarray(source)
        if img.dtype != np.uint8:
            raise TypeError("source must be 8-bit")
        if img.ndim == 2:
            img = img[:, :, None]
        if img.ndim != 3:
            raise ValueError("source must be HxW or HxWxC")
        if not img.flags.c_contiguous:
            img = np.ascontiguousarray(img)
        return img

    def _run(self, fn, source, threshold, class_ids, mask, cap):
        img = self._img(source)
        h, w, c = img.shape
        m = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            if mask.shape != (h, w):  # CV_Assert(mask.empty() || mask.size() == source.size())  :1086
                raise _capi.SbmError(-1, "mask.size() != source.size()")
            m = mask.ctypes.data_as(C.c_void_p)
        arr, n, _keep = self._class_args(class_ids)
        while True:
            out = np.zeros(cap, MATCH_REC)
            n_out = C.c_int()
            rc = fn(self._h, img.ctypes.data_as(C.c_void_p), w, h, c, 0, m, C.c_float(threshold), arr, n,
                    out.ctypes.data_as(C.c_void_p), cap, C.byref(n_out))
            if rc == -3 and n_out.value > cap:
                cap = n_out.value
                continue
            check(rc)
            return out[: n_out.value].copy()

    def match_records(self, source, threshold, class_ids=(), masks=None, cap=4096):
        """Final list as a structured array (MATCH_REC; class_idx indexes classIds())."""
        return self._run(self._L.sbm_match, source, threshold, class_ids, masks, cap)

    def match(self, source, threshold, class_ids=(), masks=None):
        """Detector::match (line2Dup.cpp:1078-1150) -> list of Match, best first."""
        recs = self.match_records(source, threshold, class_ids, masks)
        return [Match(int(r["x"]), int(r["y"]), float(r["similarity"]), self._slots[int(r["class_idx"])], int(r["template_id"]))
                for r in recs]

    def match_candidates(self, source, threshold, class_ids=(), masks=None, cap=4096):
        """Candidates before sort/unique, in the reference's serial order (line2Dup.cpp:1124-1140)."""
        return self._run(self._L.sbm_match_candidates, source, threshold, class_ids, masks, cap)

    def match_batch(self, frames, threshold, class_ids=(), masks=None, cap=None):
        """Detector::match over a stream of equally shaped frames (sbm_match_batch): several frames are kept in
        flight so that host->device copies overlap the kernels.  frames: sequence of HxW[xC] uint8 arrays (or one
        [N,H,W,C] array; pinned memory keeps the copies asynchronous).  Returns a list of MATCH_REC arrays, one
        per frame, each identical to match_records(frame)."""
        return self._batch(self._L.sbm_match_batch, frames, threshold, class_ids, masks, cap, len(frames))

    def _batch(self, fn, frames, threshold, class_ids, masks, cap, lists):
        if isinstance(frames, np.ndarray) and frames.ndim in (3, 4) and frames.dtype == np.uint8 and frames.flags.c_contiguous:
            # one [N, H, W(, C)] block: the frame addresses are an arithmetic sequence, no per-frame work on the host
            n = frames.shape[0]
            if n == 0:
                return []
            h, w = frames.shape[1:3]
            c = frames.shape[3] if frames.ndim == 4 else 1
            imgs = frames  # keeps the block alive over the call
            addr = frames.ctypes.data + np.arange(n, dtype=np.uint64) * np.uint64(h * w * c)
            ptrs = (C.c_void_p * n).from_buffer(addr)
        else:
            imgs = [self._img(f) for f in frames]
            n = len(imgs)
            if n == 0:
                return []
            h, w, c = imgs[0].shape
            for f in imgs:
                if f.shape != (h, w, c):
                    raise ValueError("all frames of a batch must have the same shape")
            ptrs = (C.c_void_p * n)(*[f.__array_interface__["data"][0] for f in imgs])  # (f.ctypes.data builds a ctypes object per frame)
        mptrs, keep = None, []
        if masks is not None:
            for m in masks:
                if m is None:
                    keep.append(None)
                else:
                    m = np.ascontiguousarray(m, dtype=np.uint8)
                    if m.shape != (h, w):
                        raise _capi.SbmError(-1, "mask.size() != source.size()")
                    keep.append(m)
            mptrs = (C.c_void_p * n)(*[(m.ctypes.data if m is not None else None) for m in keep])
        arr, ncls, _keep = self._class_args(class_ids)
        cap = cap or 256 * lists
        begin = (C.c_int * (lists + 1))()
        while True:
            out = np.empty(cap, MATCH_REC)  # the library writes begin[lists] records; nothing behind them is read
            rc = fn(self._h, ptrs, mptrs, n, w, h, c, 0, C.c_float(threshold), arr, ncls, out.ctypes.data_as(C.c_void_p), cap, begin)
            if rc == -3 and begin[lists] > cap:
                cap = begin[lists]
                continue
            check(rc)
            return self._split(out, begin, lists)

    @staticmethod
    def _split(out, begin, lists):
        """packed records + list boundaries -> one array per list (views of one trimmed copy: a ctypes int per access and a
        numpy copy per list would cost more than the C side spends ordering a short list)"""
        b = np.frombuffer(begin, dtype=np.int32, count=lists + 1).tolist()
        # (copied as bytes: numpy copies a structured array field by field, 25 ns per record against a memcpy)
        out = out.view(np.uint8)[: b[lists] * out.itemsize].copy().view(out.dtype)
        return [out[b[i] : b[i + 1]] for i in range(lists)]

    # ---- device-resident path (bench `value`, multi-GPU) ----------------------------------------
    def match_batch_device(self, d_imgs_ptr, frame_stride, n_frames, w, h, channels, threshold, class_ids, d_records, cap_per_frame,
                           d_counts, stream=0):
        arr, n, _keep = self._class_args(class_ids)
        check(self._L.sbm_match_batch_device(self._h, C.c_void_p(d_imgs_ptr), frame_stride, n_frames, w, h, channels,
                                             C.c_float(threshold), arr, n, C.c_void_p(d_records), cap_per_frame,
                                             C.c_void_p(d_counts), C.c_void_p(stream) if stream else None))

    def quantized_orientations(self, source, level=0):
        """ColorGradientPyramid::update at pyramid level `level` (line2Dup.cpp:406-444) on the GPU
        -> (magnitude f32, angle_ori f32 degrees, angle u8 one-hot), each [h >> level][w >> level]."""
        img = self._img(source)
        h, w, c = img.shape
        lw, lh = w, h
        for _ in range(level):
            lw, lh = lw // 2, lh // 2
        mag = np.zeros((lh, lw), np.float32)
        ang = np.zeros((lh, lw), np.float32)
        q = np.zeros((lh, lw), np.uint8)
        wo, ho = C.c_int(), C.c_int()
        check(self._L.sbm_quantized_orientations(self._h, img.ctypes.data_as(C.c_void_p), w, h, c, 0, level,
                                                 mag.ctypes.data_as(C.c_void_p), ang.ctypes.data_as(C.c_void_p),
                                                 q.ctypes.data_as(C.c_void_p), None, C.byref(wo), C.byref(ho)))
        assert (wo.value, ho.value) == (lw, lh)
        return mag, ang, q

    def time_kernels(self, threshold, iters=20):
        L = self.pyramid_levels
        ms = np.zeros(4 * L + 1, np.float32)
        check(self._L.sbm_time_kernels(self._h, C.c_float(threshold), iters, ms.ctypes.data_as(C.c_void_p), ms.size))
        out = {}
        for l in range(L):
            if l:
                out["pyrdown_L%d" % l] = float(ms[3 * l])
            out["orient_L%d" % l] = float(ms[3 * l + 1])
            out["spread_linearize_L%d" % l] = float(ms[3 * l + 2])
        out["similarity"] = float(ms[3 * L])
        if L > 1:
            out["refine_L%d" % (L - 2)] = float(ms[3 * L + 1])
        out["frame_graph"] = float(ms[4 * L])
        return out

    def time_similarity(self, threshold, iters=20):
        ms = C.c_float()
        check(self._L.sbm_time_similarity(self._h, C.c_float(threshold), iters, C.byref(ms)))
        return ms.value

    def time_l2_read(self, nbytes=2 << 20, iters=20):
        """GB/s of 16-byte L1-bypassing loads over an L2-resident buffer (roofline denominator for k_similarity)"""
        g = C.c_float()
        check(self._L.sbm_time_l2_read(self._h, nbytes, iters, C.byref(g)))
        return g.value

    def match_device(self, d_img_ptr, w, h, channels, threshold, class_ids=(), d_records=0, cap=0, d_count=0, stream=0,
                     d_mask=0):
        arr, n, _keep = self._class_args(class_ids)
        check(self._L.sbm_match_device(self._h, C.c_void_p(d_img_ptr), w, h, channels, C.c_void_p(d_mask) if d_mask else None,
                                       C.c_float(threshold), arr, n, C.c_void_p(d_records) if d_records else None, cap,
                                       C.c_void_p(d_count) if d_count else None, C.c_void_p(stream) if stream else None))

    def device_results(self):
        r, cnt, cap = C.c_void_p(), C.c_void_p(), C.c_int()
        check(self._L.sbm_device_results(self._h, C.byref(r), C.byref(cnt), C.byref(cap)))
        return r.value, cnt.value, cap.value

    def finalize_records(self, recs, class_ids=(), finalize=True):
        recs = np.ascontiguousarray(recs, dtype=DEVICE_REC)
        arr, n, _keep = self._class_args(class_ids)
        cap = max(1, recs.shape[0])
        out = np.zeros(cap, MATCH_REC)
        n_out = C.c_int()
        check(self._L.sbm_finalize_records(self._h, recs.ctypes.data_as(C.c_void_p), recs.shape[0], arr, n, int(finalize),
                                           out.ctypes.data_as(C.c_void_p), cap, C.byref(n_out)))
        return out[: n_out.value].copy()

    def set_candidate_capacity(self, records):
        """records each pipeline stage may hold per frame inside the context (sbm_set_candidate_capacity)"""
        check(self._L.sbm_set_candidate_capacity(self._h, int(records)))

    # ---- multi-GPU group: one Detector per GPU, NCCL behind the C ABI (sbm_group_*) ---------------
    @staticmethod
    def group_unique_id():
        """128 bytes from rank 0 (ncclGetUniqueId) that every rank hands to group_create"""
        _capi.prefer_bundled_nccl()
        buf = C.create_string_buffer(_capi.GROUP_ID_BYTES)
        check(_capi.lib().sbm_group_unique_id(buf))
        return buf.raw

    def group_create(self, n_ranks, rank, unique_id, mode=_capi.GROUP_TEMPLATES):
        """join a group of n_ranks detectors (collective).  Template modes: keep calling setClassTemplates with the
        FULL classes on every rank, the context keeps this rank's slice."""
        assert len(unique_id) == _capi.GROUP_ID_BYTES
        _capi.prefer_bundled_nccl()
        check(self._L.sbm_clear_templates(self._h))
        self._dirty = True
        check(self._L.sbm_group_create(self._h, int(n_ranks), int(rank), C.c_char_p(unique_id), int(mode)))
        self._group = (int(n_ranks), int(rank), int(mode))

    def group_destroy(self):
        check(self._L.sbm_group_destroy(self._h))
        self._group = None
        self._dirty = True

    def group_info(self):
        return getattr(self, "_group", None)

    def group_match_records(self, source, threshold, class_ids=(), masks=None, cap=4096):
        """Detector::match with the templates sharded over the group: the unsharded final list, on every rank"""
        return self._run(self._L.sbm_group_match, source, threshold, class_ids, masks, cap)

    def group_match_batch(self, frames, threshold, class_ids=(), masks=None, cap=None):
        """sbm_group_match_batch: template sharding -> one list per frame; frame sharding -> the lists of all ranks'
        frames, rank-major (every rank passes the same number of frames)"""
        n_ranks, _rank, mode = self._group
        lists = len(frames) * (n_ranks if mode == _capi.GROUP_FRAMES else 1)
        return self._batch(self._L.sbm_group_match_batch, frames, threshold, class_ids, masks, cap, lists)

    @staticmethod
    def group_block_bytes(cap_per_frame):
        return int(_capi.lib().sbm_group_block_bytes(int(cap_per_frame)))

    def group_match_batch_device(self, d_imgs_ptr, frame_stride, n_frames, w, h, channels, threshold, class_ids, d_gathered, cap_per_frame,
                                 stream=0):
        """device-resident, asynchronous on `stream`: d_gathered = uint8 [n_ranks][n_frames][group_block_bytes(cap)]"""
        arr, n, _keep = self._class_args(class_ids)
        check(self._L.sbm_group_match_batch_device(self._h, C.c_void_p(d_imgs_ptr), frame_stride, n_frames, w, h, channels,
                                                   C.c_float(threshold), arr, n, C.c_void_p(d_gathered), cap_per_frame,
                                                   C.c_void_p(stream) if stream else None))

    def group_finalize(self, h_gathered, n_frames, cap_per_frame, class_ids=(), finalize=True, cap=None):
        """host copy of d_gathered -> list of MATCH_REC arrays (laid out like group_match_batch's result)"""
        n_ranks, _rank, mode = self._group
        lists = n_frames * (n_ranks if mode == _capi.GROUP_FRAMES else 1)
        h = np.ascontiguousarray(h_gathered, dtype=np.uint8).ravel()
        assert h.size == n_ranks * n_frames * self.group_block_bytes(cap_per_frame)
        arr, n, _keep = self._class_args(class_ids)
        cap = cap or max(256 * lists, 1)
        begin = (C.c_int * (lists + 1))()
        while True:
            out = np.zeros(cap, MATCH_REC)
            rc = self._L.sbm_group_finalize(self._h, h.ctypes.data_as(C.c_void_p), n_frames, cap_per_frame, arr, n, int(bool(finalize)),
                                            out.ctypes.data_as(C.c_void_p), cap, begin)
            if rc == -3 and begin[lists] > cap and "output holds" in _capi.lib().sbm_last_error().decode():
                cap = begin[lists]
                continue
            check(rc)
            return self._split(out, begin, lists)

    # ---- taps for parity tests ------------------------------------------------------------------
    def set_debug(self, on=True):
        check(self._L.sbm_set_debug(self._h, int(on)))

    def set_timing(self, on=True):
        check(self._L.sbm_set_timing(self._h, int(on)))

    def intermediate(self, level, kind):
        need = C.c_size_t()
        check(self._L.sbm_get_intermediate(self._h, level, kind, None, 0, C.byref(need)))
        buf = np.zeros(need.value, np.uint8)
        check(self._L.sbm_get_intermediate(self._h, level, kind, buf.ctypes.data_as(C.c_void_p), need.value, None))
        return buf

    def similarity_map(self, class_id, template_id, shape):
        self._sync_templates()
        out = np.zeros(shape, np.uint16)
        check(self._L.sbm_similarity_map(self._h, self._slots.index(class_id), template_id, out.ctypes.data_as(C.c_void_p),
                                         out.nbytes))
        return out

    def similarity_local(self, class_id, template_id, level, x, y):
        self._sync_templates()
        out = np.zeros((16, 16), np.uint16)
        check(self._L.sbm_similarity_local(self._h, self._slots.index(class_id), template_id, level, x, y,
                                           out.ctypes.data_as(C.c_void_p)))
        return out

    def last_stats(self):
        s = np.zeros(8, np.uint64)
        check(self._L.sbm_last_stats(self._h, s.ctypes.data_as(C.c_void_p)))
        return dict(coarse_evals=int(s[0]), coarse_candidates=int(s[1]), local_evals=int(s[2]), records=int(s[3]),
                    gpu_launches=int(s[4]), coarse_nibbles=int(int(s[5]) == 1), coarse_layout=("byte", "nib", "bits")[int(s[5])],
                    refine_tiled=int(s[6]))

    def last_timing(self):
        t = np.zeros(4, np.float32)
        check(self._L.sbm_last_timing(self._h, t.ctypes.data_as(C.c_void_p)))
        return dict(total_ms=float(t[0]), front_ms=float(t[1]), similarity_ms=float(t[2]), refine_ms=float(t[3]))
