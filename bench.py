#!/usr/bin/env python
"""bench.py -- Detector::match hot path on B200: template-pixel evaluations/s and ms/frame.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--frames B] [--skip c4,c5,cpu]

Headline workload (BASELINE.json metric, config M1 of BASELINE.md): 1024x1024x3 frames x 360 rotation templates
(test/case1 templates 0..359, 131/71 features at the two pyramid levels, T={4,8}, threshold 90).  One STEP = one pass of
the whole hot path (orientations -> spread -> response maps -> linear memories -> similarity -> thresholded candidates
-> local refinement) over a batch of B distinct synthetic frames per GPU (default 256 frames = 768 MiB of input per
rank, larger than the 126 MB L2, so no frame is cache-warm; 20 steps last ~0.1 s).

  value   template-pixel evaluations/s, frames already resident in HBM, results left in HBM, timed with CUDA events on
          the launching stream (max over ranks); `step_ms` gives min / median / max over the K steps.
  e2e     same metric through the public API with HOST buffers: Detector.match_batch(frames) from pinned host memory to
          the final host lists (H2D + pipeline + D2H + sort/unique for every frame; several frames in flight).
  N > 1   M1 shards FRAMES (rank r owns its own B frames, all 360 templates: weak scaling); the per-frame result blocks
          are exchanged at the end of the step with ONE ncclAllGather issued inside the C ABI (sbm_group_match_batch_
          device) -- the only exchange the path has.  torch.distributed carries the NCCL id, the barrier and the
          max-over-ranks of the timings, nothing of the data path.
  c4_template_sharded   BASELINE.json config 4 (M4): one 4096x4096 frame x 3600 templates (0.1 degree x 3 scales, trained
          with num_features 8191 through shapeInfo_producer / addTemplate / addTemplate_rotate), TEMPLATES sharded over
          the N GPUs (strong scaling), merged list checked on every run against the unsharded list.
  c5_frame_sharded      BASELINE.json config 5 (M5): 256 frames of 2048x2048 x 720 templates, FRAMES sharded, frames/s
          device-resident and from pinned host memory.
  --impl reference   the reference's own CPU code (oracle/_ref = /root/reference/line2Dup.cpp compiled verbatim, OpenMP
          over all host threads) on a bounded sample of the same frames.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import threading
import time

# eight frame graphs of two branches each in flight: more hardware queues than the default 8 (before CUDA is initialised;
# the library asks for the same when it is loaded first)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "template_pixel_evals_per_s"
UNIT = "evals/s"
THRESHOLD = 90.0
W = H = 1024
C = 3
N_TEMPLATES = 360
REC_CAP = 512  # per-frame record capacity of the device-resident path = the allgather payload (20 B each); M1 frames
# produce ~100 records; the block headers are checked after the timed region, an overflow aborts the run
_JSON_FD = None  # the original stdout when fd 1 has been pointed at stderr (N > 1)
M1_NAME = "M1: 1024x1024x3 frames x 360 templates (case1 ids 0..359, 131/71 features, T={4,8}, thr 90)"


def load_workload():
    from _fixtures import load_case, load_templates

    case = load_case("m1")
    cid, tdesc, feats, nf = load_templates("case1[:360]")
    return case["img"], cid, tdesc, feats, nf


def make_frames(base, n, seed0):
    """n distinct frames derived from the M1 frame: cyclic shifts (multiples of 16 keep every true match
    on the coarse grid) + a per-frame sprinkle of low-amplitude noise."""
    out = np.empty((n,) + base.shape, np.uint8)
    for i in range(n):
        rng = np.random.default_rng(seed0 + i)
        dy, dx = int(rng.integers(0, 64)) * 16, int(rng.integers(0, 64)) * 16
        f = np.roll(base, (dy, dx), axis=(0, 1)).astype(np.int16)
        f += rng.integers(-2, 3, f.shape, dtype=np.int16)
        out[i] = np.clip(f, 0, 255).astype(np.uint8)
    return out


def coarse_evals_per_frame(tdesc, feats, T=8, w=512, h=512):
    """sum over templates of (features inside the coarse image) * template_positions (line2Dup.cpp:813-857)."""
    Wc, Hc = w // T, h // T
    total = 0
    fo = 0
    for t in range(tdesc.shape[0]):
        for l in range(tdesc.shape[1]):
            nf = int(tdesc[t, l, 4])
            if l == tdesc.shape[1] - 1:
                f = feats[fo : fo + nf]
                nv = int(((f[:, 0] >= 0) & (f[:, 0] < w) & (f[:, 1] >= 0) & (f[:, 1] < h)).sum())
                wf = (int(tdesc[t, l, 0]) - 1) // T + 1
                hf = (int(tdesc[t, l, 1]) - 1) // T + 1
                tp = (Hc - hf) * Wc + (Wc - wf) + 1
                total += nv * max(tp, 0)
            fo += nf
    return total


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def best_threads(det, frame, cid, cores):
    """The reference arm gets the OpenMP thread count that serves it best on this host (all cores, half, or one:
    on hosts with a CPU quota below the visible core count, spinning OpenMP teams lose to a single thread)."""
    best, best_t = cores, None
    for n in sorted({cores, max(1, cores // 2), 1}, reverse=True):
        det.match(frame, THRESHOLD, [cid], omp_threads=n)
        t0 = time.perf_counter()
        for _ in range(2):
            det.match(frame, THRESHOLD, [cid], omp_threads=n)
        t = time.perf_counter() - t0
        if best_t is None or t < best_t:
            best, best_t = n, t
    return best


def cv2_front_end_ms(frame, iters=5):
    """the OpenCV image ops of the front end (GaussianBlur, Sobel x2, magnitude, phase, convertTo,
    pyrDown; both pyramid levels) timed through REAL OpenCV (Python cv2), next to the restated ops the compiled reference
    runs.  Only the OpenCV calls: the reference's own loops (hysteresis, spread, response maps, linearize) are not here."""
    try:
        import cv2
    except Exception as ex:
        return {"unavailable": str(ex)}

    def level(img):
        sm = cv2.GaussianBlur(img, (7, 7), 0, borderType=cv2.BORDER_REPLICATE)
        dx = cv2.Sobel(sm, cv2.CV_16S, 1, 0, ksize=3, borderType=cv2.BORDER_REPLICATE)
        dy = cv2.Sobel(sm, cv2.CV_16S, 0, 1, ksize=3, borderType=cv2.BORDER_REPLICATE)
        fx, fy = dx[..., 0].astype(np.float32), dy[..., 0].astype(np.float32)  # one channel's worth of the float stage
        cv2.magnitude(fx, fy)
        ang = cv2.phase(fx, fy, angleInDegrees=True)
        cv2.convertScaleAbs(ang, alpha=16.0 / 360.0)

    def run():
        level(frame)
        level(cv2.pyrDown(frame, dstsize=(frame.shape[1] // 2, frame.shape[0] // 2)))

    run()
    t0 = time.perf_counter()
    for _ in range(iters):
        run()
    return {"ms_per_frame": (time.perf_counter() - t0) / iters * 1e3, "threads": int(cv2.getNumThreads()), "cv2": cv2.__version__,
            "ops": "GaussianBlur 7x7, Sobel dx + dy (3 ch), magnitude, phase, convertScaleAbs, pyrDown; levels 0 and 1"}


def reference_arm(args, rank, world):
    """The reference's own CPU implementation of the path, all host threads (rank 0 only)."""
    if rank != 0:
        return
    from oracle import ref

    base, cid, tdesc, feats, nf = load_workload()
    cores = os.cpu_count() or 1
    det = ref.RefDetector(nf, (4, 8))
    det.add_class(cid, tdesc, feats)
    sample = max(1, min(args.ref_frames, args.frames))
    frames = make_frames(base, sample, 1000)
    # evaluations of the sample: coarse (closed form) + local (256 * nf per coarse candidate), counted by
    # the reference's own similarity maps (oracle/ref_driver.cpp ref_eval_counts)
    evals = 0
    for f in frames:
        ce, _cc, le = det.eval_counts(f, THRESHOLD, [cid])
        assert ce == coarse_evals_per_frame(tdesc, feats)
        evals += ce + le
    cores = best_threads(det, frames[0], cid, cores)
    for _ in range(args.warmup):
        det.match(frames[0], THRESHOLD, [cid], omp_threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for f in frames:
            det.match(f, THRESHOLD, [cid], omp_threads=cores)
    dt = (time.perf_counter() - t0) / args.steps
    value = evals / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "ms_per_frame": dt * 1e3 / sample, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": M1_NAME, "frames_per_step": sample, "evals_per_frame": evals / sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "%d frames/step x %d steps, Detector::match verbatim (oracle/_ref, -O3 -mavx2 -mfma -fopenmp, "
                                   "OpenCV ops restated in oracle/shim, row-parallel + auto-vectorised)" % (sample, args.steps),
                         "real_opencv_front_end": cv2_front_end_ms(frames[0])},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa(local_rank):
    """Run this rank's host threads (and with them its pinned allocations, first-touch) on the NUMA node its GPU hangs off.
    Best effort; returns what was done for the JSON line."""
    try:
        import torch

        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return {"numa_node": node, "bound": False, "why": "no NUMA information for the GPU"}
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = sorted(set(cpus) & allowed)
        if not use:
            return {"numa_node": node, "bound": False, "why": "node's CPUs are outside this process's affinity mask"}
        os.sched_setaffinity(0, use)
        return {"numa_node": node, "bound": True, "cpus": len(use)}
    except Exception as ex:
        return {"bound": False, "why": "%s: %s" % (type(ex).__name__, ex)}


def same_records(a, b):
    return (len(a) == len(b) and np.array_equal(a["x"], b["x"]) and np.array_equal(a["y"], b["y"]) and
            np.array_equal(a["template_id"], b["template_id"]) and np.array_equal(a["similarity"].view(np.uint32), b["similarity"].view(np.uint32)))


def latest_profile_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the newest committed
    profiles/*_traffic.json that names it ({"commit": ..., "kernels": {name: {"dram_bytes_per_launch": ...}}});
    bench.py itself never runs under a profiler."""
    best = None
    for p in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json"))):
        try:
            d = json.load(open(p))
            if kernel in d.get("kernels", {}):
                best = (d["kernels"][kernel].get("dram_bytes_per_launch"), os.path.relpath(p, ROOT), d.get("commit"))
        except Exception:
            pass
    return best


class Ranks:
    """torch.distributed as plumbing: barrier, max / sum over ranks, the NCCL ids of the detectors' own groups."""

    def __init__(self, world, dev):
        self.world, self.dev = world, dev
        if world > 1:
            import torch.distributed as dist

            self.dist = dist

    def barrier(self):
        import torch

        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def reduce(self, vals, op="max"):
        import torch

        t = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return [float(v) for v in t]

    def gather_objects(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out

    def unique_id(self):
        from shape_based_matching_b200 import Detector

        if self.world == 1:
            return Detector.group_unique_id()
        box = [Detector.group_unique_id() if self.dist.get_rank() == 0 else None]
        self.dist.broadcast_object_list(box, src=0)
        return box[0]


def timed_steps(ranks, stream, step, steps, warmup):
    """W untimed + K timed calls of step(); CUDA events on the launching stream, barrier + synchronize on both sides;
    -> (total ms of the K steps, [ms of every step])"""
    import torch

    for _ in range(warmup):
        step()
    ranks.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ranks.barrier()
    ev[0].record(stream)
    for k in range(steps):
        step()
        ev[k + 1].record(stream)
    ranks.barrier()
    per = [ev[k].elapsed_time(ev[k + 1]) for k in range(steps)]
    return ev[0].elapsed_time(ev[steps]), per


def c4_leg(args, ranks, rank, world, local_rank, stream):
    """BASELINE.json config 4 / BASELINE.md M4, templates sharded over the ranks (the template loop of
    matchClass, /root/reference/line2Dup.cpp:1166-1170, + the reduction :1295)."""
    import torch

    from shape_based_matching_b200 import Detector
    from shape_based_matching_b200 import _capi
    from shape_based_matching_b200 import workloads as wl

    t_setup = time.perf_counter()
    M = wl.M4
    tdesc, feats, infos = wl.m4_templates(device=local_rank)
    frame = wl.m4_frame()
    hw = M["hw"]
    thr = M["threshold"]
    uid = ranks.unique_id()
    det = Detector(M["num_features"], (4, 8), 30.0, 60.0, device=local_rank)
    det.group_create(world, rank, uid, _capi.GROUP_TEMPLATES)
    det.setClassTemplates(M["class_id"], tdesc, feats)
    cid = [M["class_id"]]
    merged = det.group_match_records(frame, thr, cid, cap=1 << 16)  # through the real NCCL exchange
    st = det.last_stats()
    # parity on every run: rank 0 also matches the whole set unsharded
    parity = None
    if rank == 0:
        plain = Detector(M["num_features"], (4, 8), 30.0, 60.0, device=local_rank)
        plain.setClassTemplates(M["class_id"], tdesc, feats)
        whole = plain.match_records(frame, thr, cid, cap=1 << 16)
        if not same_records(merged, whole):
            raise SystemExit("bench.py: C4 merged template-sharded list (%d) disagrees with the unsharded list (%d)" % (len(merged), len(whole)))
        pst = plain.last_stats()
        total_evals = pst["coarse_evals"] + pst["local_evals"]
        planted = [(12.3, 1.0), (77.7, 0.9), (101.0, 1.1)]
        found = []
        for ang, sc in planted:
            ok = [r for r in whole if abs(float(infos[int(r["template_id"])][0]) - ang) < 0.35 and abs(float(infos[int(r["template_id"])][1]) - sc) < 0.01]
            found.append(len(ok) > 0)
        parity = {"merged_equals_unsharded": True, "matches": int(len(whole)), "planted_instances_found": found}
        plain.close()
    else:
        total_evals = 0
    total_evals = ranks.reduce([float(total_evals)], "max")[0]
    sig = float(len(merged)) + (float(np.sum(merged["x"].astype(np.int64) * 7 + merged["template_id"])) if len(merged) else 0.0)
    lo, hi = ranks.reduce([-sig], "max")[0], ranks.reduce([sig], "max")[0]
    if -lo != hi:
        raise SystemExit("bench.py: the ranks hold different merged C4 lists")
    # device-resident timing: n_f copies of the frame per step
    n_f = 2
    d_frames = torch.from_numpy(np.stack([frame] * n_f)).to(stream.device)
    cap = 4096
    blk = Detector.group_block_bytes(cap)
    d_g = torch.zeros(world * n_f * blk, dtype=torch.uint8, device=stream.device)

    def step():
        det.group_match_batch_device(d_frames.data_ptr(), hw * hw, n_f, hw, hw, 1, thr, cid, d_g.data_ptr(), cap, stream.cuda_stream)

    steps = max(2, min(args.steps, 4))
    ms, per = timed_steps(ranks, stream, step, steps, 1)
    lists = det.group_finalize(d_g.cpu().numpy(), n_f, cap, cid)
    if not same_records(lists[0], merged) or not same_records(lists[1], merged):
        raise SystemExit("bench.py: C4 device-resident group path disagrees with the host group path")
    ms = ranks.reduce([ms], "max")[0]
    ms_frame = ms / steps / n_f
    # per-rank split (warm kernels of this rank's shard)
    det.match_records(frame, thr, cid, cap=1 << 16)  # plain call: fills slot 0 for the kernel timers
    km = det.time_kernels(thr, iters=2)
    split = {"front_end_ms": sum(v for k, v in km.items() if k.startswith(("orient", "pyrdown", "spread"))), "similarity_ms": km["similarity"],
             "refine_ms": km.get("refine_L0", 0.0), "templates": det.numTemplates(), "coarse_evals": det.last_stats()["coarse_evals"]}
    splits = ranks.gather_objects(split)
    # end to end from pinned host memory through the group call
    pinned = torch.from_numpy(np.stack([frame] * n_f)).pin_memory().numpy()
    host_frames = [pinned[i] for i in range(n_f)]
    det.group_match_batch(host_frames, thr, cid, cap=1 << 16)
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        det.group_match_batch(host_frames, thr, cid, cap=1 << 16)
    torch.cuda.synchronize()
    e2e = ranks.reduce([(time.perf_counter() - t0) / steps / n_f], "max")[0]
    det.close()
    del st
    return {"workload": "M4: 4096x4096x1 frame (seed 4096) x %d templates (0.1 deg x scales 0.9/1.0/1.1; num_features request 8191, lowered 5%% at a "
                        "time until Detector::match's 8192 assert passes), thr %.0f" % (tdesc.shape[0], thr),
            "features_L0_L1_per_scale": [[int(v) for v in tdesc[i, :, 4]] for i in (0, tdesc.shape[0] // 3, 2 * (tdesc.shape[0] // 3))],
            "sharding": "templates, interleaved ids (rank r holds r, r+N, ...); every rank builds the frame's linear memories; one ncclAllGather of "
                        "result blocks per step inside sbm_group_match_batch_device", "scaling": "strong",
            "ms_per_frame": ms_frame, "evals_per_frame": total_evals, "evals_per_s": total_evals / (ms_frame * 1e-3),
            "frames_per_step": n_f, "steps": steps, "step_ms": {"min": min(per), "median": float(np.median(per)), "max": max(per)},
            "e2e_ms_per_frame": e2e * 1e3, "exchange_bytes_per_rank_per_step": n_f * blk, "per_rank": splits, "parity": parity,
            "setup_s": time.perf_counter() - t_setup}


def c5_leg(args, ranks, rank, world, local_rank, stream):
    """BASELINE.json config 5 / BASELINE.md M5, frames sharded over the ranks (the caller's frame loop,
    /root/reference/test_jabil.cpp:314-374)."""
    import torch

    from shape_based_matching_b200 import Detector
    from shape_based_matching_b200 import _capi
    from shape_based_matching_b200 import workloads as wl

    t_setup = time.perf_counter()
    M = wl.M5
    tdesc, feats, _infos = wl.m5_templates(device=local_rank)
    hw, thr = M["hw"], M["threshold"]
    total = args.c5_frames
    per = total // world
    frames = np.stack([wl.m5_frame(rank * per + i) for i in range(per)])
    uid = ranks.unique_id()
    det = Detector(M["num_features"], (4, 8), 30.0, 60.0, device=local_rank)
    det.group_create(world, rank, uid, _capi.GROUP_FRAMES)
    det.setClassTemplates(M["class_id"], tdesc, feats)
    cid = [M["class_id"]]
    evals = 0
    own = []
    for i in range(per):
        own.append(det.match_records(frames[i], thr, cid))
        st = det.last_stats()
        evals += st["coarse_evals"] + st["local_evals"]
    cap = 1024
    blk = Detector.group_block_bytes(cap)
    d_frames = torch.from_numpy(frames).to(stream.device)
    d_g = torch.zeros(world * per * blk, dtype=torch.uint8, device=stream.device)

    def step():
        det.group_match_batch_device(d_frames.data_ptr(), hw * hw, per, hw, hw, 1, thr, cid, d_g.data_ptr(), cap, stream.cuda_stream)

    steps = max(2, min(args.steps, 5))
    ms, per_ms = timed_steps(ranks, stream, step, steps, 2)
    lists = det.group_finalize(d_g.cpu().numpy(), per, cap, cid)
    for i in range(per):  # this rank's frames, as every rank received them through the exchange
        if not same_records(lists[rank * per + i], own[i]):
            raise SystemExit("bench.py: C5 frame %d of rank %d changed in the exchange" % (i, rank))
    parity = {"gathered_equals_single_frame_calls": True, "records_first_frames": [int(len(l)) for l in lists[:4]]}
    if rank == 0 and not args.no_cpu_baseline:
        try:
            from oracle import ref

            if ref.available():
                rd = ref.RefDetector(M["num_features"], (4, 8))
                rd.add_class(M["class_id"], tdesc, feats)
                want = rd.match(frames[0], thr, cid, omp_threads=os.cpu_count() or 1)
                # the reference's multi-thread list is a permutation-stable set here only at one thread; compare as sets
                want1 = rd.match(frames[0], thr, cid, omp_threads=1)
                parity["frame0_equals_reference"] = bool(same_records(lists[0], want1))
                parity["frame0_reference_records"] = int(len(want1))
                del want
        except Exception as ex:
            parity["frame0_equals_reference"] = "unavailable: %s" % ex
    ms = ranks.reduce([ms], "max")[0]
    evals_all = ranks.reduce([float(evals)], "sum")[0]
    fps = per * world / (ms / steps * 1e-3)
    pinned = torch.from_numpy(frames).pin_memory().numpy()
    host_frames = [pinned[i] for i in range(per)]
    det.group_match_batch(host_frames, thr, cid)
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        det.group_match_batch(host_frames, thr, cid)
    torch.cuda.synchronize()
    e2e_s = ranks.reduce([(time.perf_counter() - t0) / steps], "max")[0]
    det.close()
    return {"workload": "M5: %d frames of 2048x2048x1 (seeds 0..%d) x %d templates (0.5 deg, %d/%d features), thr %.0f" % (
                per * world, per * world - 1, tdesc.shape[0], int(tdesc[0, 0, 4]), int(tdesc[0, 1, 4]), thr),
            "sharding": "frames (rank r owns frames r*%d .. ; all templates on every rank); one ncclAllGather of the per-frame result blocks per pass "
                        "inside sbm_group_match_batch_device" % per, "scaling": "strong",
            "frames_per_s": fps, "ms_per_frame": ms / steps / (per * world), "evals_per_s": evals_all / (ms / steps * 1e-3),
            "frames_per_s_from_pinned_host": per * world / e2e_s, "frames_per_rank": per, "steps": steps,
            "step_ms": {"min": min(per_ms), "median": float(np.median(per_ms)), "max": max(per_ms)},
            "h2d_bytes_per_pass_per_rank": int(frames.nbytes), "exchange_bytes_per_rank_per_pass": per * blk, "parity": parity,
            "setup_s": time.perf_counter() - t_setup}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=256, help="M1 frames per step per rank")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the CPU reference arm")
    ap.add_argument("--cpu-baseline-frames", type=int, default=8)
    ap.add_argument("--c5-frames", type=int, default=256, help="frames of configuration 5 (all ranks together)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--skip", default="", help="comma list of legs to skip: c4, c5, cpu")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    skip = set(s for s in args.skip.split(",") if s)
    if "cpu" in skip:
        args.no_cpu_baseline = True

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from shape_based_matching_b200 import _capi
    from shape_based_matching_b200 import Detector

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa(local_rank)
    if world > 1:
        # NCCL prints its version banner (NCCL_DEBUG=VERSION/INFO on the box) with the C library on fd 1: point fd 1 at
        # stderr while the process group lives, so that stdout carries the one JSON line and nothing else
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    ranks = Ranks(world, dev)

    base, cid, tdesc, feats, nf = load_workload()
    B = args.frames
    frames = make_frames(base, B, 1000 + rank * B)
    det = Detector(nf, (4, 8), 30.0, 60.0, device=local_rank)
    if world > 1:  # M1 shards frames: the detector's own NCCL communicator carries the exchange (sbm_group_*)
        det.group_create(world, rank, ranks.unique_id(), _capi.GROUP_FRAMES)
    det.setClassTemplates(cid, tdesc, feats)

    # ---- parity guard: the frame the fixture was generated from must still give the golden list ----
    from _fixtures import load_case

    gold = load_case("m1")
    got = det.match_records(gold["img"], THRESHOLD, [cid])
    if not same_records(got, gold["matches"]):
        raise SystemExit("bench.py: CUDA path disagrees with the golden reference list; refusing to time it")

    d_frames = torch.from_numpy(frames).to(dev)  # [B, H, W, C] resident in HBM
    # a real (non-default) stream: the C ABI treats a NULL stream as "the detector's own stream", and the timing
    # events must sit on the stream the kernels are launched on
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    assert sptr != 0
    rec_bytes = _capi.DEVICE_REC.itemsize
    blk = Detector.group_block_bytes(REC_CAP)
    d_recs = torch.zeros((B, REC_CAP * rec_bytes), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros((B, 2), dtype=torch.int32, device=dev)
    d_gath = torch.zeros(world * B * blk, dtype=torch.uint8, device=dev) if world > 1 else None
    frame_bytes = W * H * C

    def step_device():
        if world > 1:  # the B frames over the frame slots, then the path's one exchange: result blocks, allgathered over NVLink
            det.group_match_batch_device(d_frames.data_ptr(), frame_bytes, B, W, H, C, THRESHOLD, [cid], d_gath.data_ptr(), REC_CAP, sptr)
        else:
            det.match_batch_device(d_frames.data_ptr(), frame_bytes, B, W, H, C, THRESHOLD, [cid], d_recs.data_ptr(), REC_CAP,
                                   d_cnt.data_ptr(), sptr)

    def step_device_serial(n):
        for i in range(n):
            det.match_device(d_frames[i].data_ptr(), W, H, C, THRESHOLD, [cid], d_records=d_recs[i].data_ptr(), cap=REC_CAP,
                             d_count=d_cnt[i].data_ptr(), stream=sptr)

    # evaluation count of one step on this rank (coarse: closed form; local: counted by the kernels)
    local_evals = 0
    singles = []
    for i in range(B):
        singles.append(det.match_records(frames[i], THRESHOLD, [cid]))
        st = det.last_stats()
        local_evals += st["local_evals"]
    launches_per_frame = st["gpu_launches"]
    coarse = st["coarse_evals"]
    assert coarse == coarse_evals_per_frame(tdesc, feats), "evaluation count mismatch"
    evals_step_rank = coarse * B + local_evals

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms, per_step = timed_steps(ranks, stream, step_device, args.steps, args.warmup)
    if world > 1:
        # every frame of every rank, as this rank received it, equals the single-frame call (own frames) / is complete
        lists = det.group_finalize(d_gath.cpu().numpy(), B, REC_CAP, [cid])
        max_recs = max(len(l) for l in lists)
        for i in range(B):
            if not same_records(lists[rank * B + i], singles[i]):
                raise SystemExit("bench.py: frame %d of rank %d changed in the exchange" % (i, rank))
    else:
        max_recs = int(d_cnt[:, 0].max().item())
        if max_recs > REC_CAP or int(d_cnt[:, 1].max().item()) != 0:
            raise SystemExit("bench.py: a frame produced %d records (REC_CAP=%d) or raised the overflow flag" % (max_recs, REC_CAP))
    ms = ranks.reduce([ms], "max")[0]
    evals_step = ranks.reduce([float(evals_step_rank)], "sum")[0]
    ms_per_step = ms / args.steps
    value = evals_step / (ms_per_step * 1e-3)

    # ---- single-stream latency per frame (frames strictly one after the other) ----
    nser = min(B, 64)
    step_device_serial(nser)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    step_device_serial(nser)
    e1.record(stream)
    torch.cuda.synchronize()
    serial_ms_per_frame = e0.elapsed_time(e1) / nser

    # ---- per-stage device times (event timing on; the events themselves cost a few us per stage) ----
    det.set_timing(True)
    kt = {"front_ms": [], "similarity_ms": [], "refine_ms": [], "total_ms": []}
    for i in range(min(B, 16)):
        det.match_device(d_frames[i].data_ptr(), W, H, C, THRESHOLD, [cid], d_records=d_recs[i].data_ptr(), cap=REC_CAP,
                         d_count=d_cnt[i].data_ptr(), stream=sptr)
        torch.cuda.synchronize()
        tm_ = det.last_timing()
        for k in kt:
            kt[k].append(tm_[k])
    det.set_timing(False)
    kmean = {k: float(np.mean(v)) for k, v in kt.items()}
    # the roofline kernel alone: back-to-back launches of the pipeline's own k_similarity grid between two events
    sim_ms = det.time_similarity(THRESHOLD, iters=50)
    kernel_ms = det.time_kernels(THRESHOLD, iters=20)
    stats = det.last_stats()
    layout = stats.get("coarse_layout", "nib" if stats.get("coarse_nibbles", 0) else "byte")
    bits_per_resp = {"byte": 8, "nib": 4, "bits": 2}[layout]
    # the coarse level's linear memories: 8 labels x (512 x 512) responses
    lm_bytes = 8 * (W // 2) * (H // 2) * bits_per_resp // 8
    l2_gbs = det.time_l2_read(lm_bytes, iters=20)

    # ---- end to end through the public API: pinned host frame -> final host list, every frame ----
    pinned = torch.from_numpy(frames).pin_memory()
    pin_np = pinned.numpy()
    n_final = 0
    frame_list = pin_np  # one [B, H, W, C] block of pinned memory (Detector.match_batch takes a list of frames as well)
    # N > 1: the group call, which also exchanges the lists (every rank ends up with the final lists of all ranks' frames)
    e2e_call = det.group_match_batch if world > 1 else det.match_batch
    e2e_call(frame_list, THRESHOLD, [cid])
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        n_final += sum(len(r) for r in e2e_call(frame_list, THRESHOLD, [cid]))
    torch.cuda.synchronize()
    e2e_s = ranks.reduce([(time.perf_counter() - t0) / args.steps], "max")[0]
    e2e_value = evals_step / e2e_s
    # latency of ONE blocking Detector.match call (H2D + pipeline + D2H + sort/unique), frames one at a time
    t0 = time.perf_counter()
    for i in range(nser):
        det.match_records(pin_np[i], THRESHOLD, [cid])
    single_call_ms = (time.perf_counter() - t0) / nser * 1e3
    d2h = (16 * 4 + 1024 * rec_bytes) * B  # counters + head of the record buffer, per frame
    if world > 1:  # the gathered result blocks of all ranks (256 records per block unless a frame needed more)
        d2h = world * B * Detector.group_block_bytes(256)
    # the floor under e2e: the same B pinned frames copied host -> device back to back and nothing else
    d_sink = torch.empty((B, H, W, C), dtype=torch.uint8, device=dev)
    h2d_best = None
    for _ in range(3):
        ranks.barrier()
        t0 = time.perf_counter()
        for i in range(B):
            d_sink[i].copy_(pinned[i], non_blocking=True)
        torch.cuda.synchronize()
        dt_ = time.perf_counter() - t0
        h2d_best = dt_ if h2d_best is None else min(h2d_best, dt_)
    h2d_us_per_frame = ranks.reduce([h2d_best / B * 1e6], "max")[0]
    del d_sink
    clocks = sampler.stop() if rank == 0 else None

    cpu_baseline = None
    if rank == 0 and not args.no_cpu_baseline:
        try:
            from oracle import ref

            cores = os.cpu_count() or 1
            rd = ref.RefDetector(nf, (4, 8))
            rd.add_class(cid, tdesc, feats)
            nsamp = max(1, min(args.cpu_baseline_frames, B))
            cores = best_threads(rd, frames[0], cid, cores)
            rd.match(frames[0], THRESHOLD, [cid], omp_threads=cores)
            t0 = time.perf_counter()
            for i in range(nsamp):
                rd.match(frames[i], THRESHOLD, [cid], omp_threads=cores)
            dt = time.perf_counter() - t0
            split = rd.time(frames[0], THRESHOLD, [cid], omp_threads=cores, iters=2)
            cpu_baseline = {"value": (evals_step_rank / B) * nsamp / dt, "unit": UNIT, "cores": cores, "kind": "reference",
                            "sample": "%d of the %d frames, Detector::match verbatim via oracle/_ref (-O3 -mavx2 -mfma "
                                      "-fopenmp; OpenCV ops restated in oracle/shim, row-parallel + auto-vectorised)" % (nsamp, B),
                            "ms_per_frame": dt / nsamp * 1e3, "ms_front": split["ms_front"], "ms_match": split["ms_match"],
                            "real_opencv_front_end": cv2_front_end_ms(frames[0])}
        except Exception as ex:  # the checker is optional on a box without oracle/_ref
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "unavailable: %s" % ex}

    # M1's detector is done: free its frames before the big configurations
    del d_frames, d_recs, d_cnt, d_gath, pinned, pin_np, frame_list
    det.close()
    torch.cuda.empty_cache()
    c4 = c5 = None
    if "c4" not in skip:
        c4 = c4_leg(args, ranks, rank, world, local_rank, stream)
        torch.cuda.empty_cache()
    if "c5" not in skip:
        c5 = c5_leg(args, ranks, rank, world, local_rank, stream)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    # dominant kernel = the coarse similarity kernel: 1 algorithmic byte per coarse evaluation
    kname = {"byte": "k_similarity", "nib": "k_similarity_nib", "bits": "k_similarity_bits"}[layout]
    sim_gbs = coarse / (sim_ms * 1e-3) / 1e9
    front_bytes = (W * H * C + 8 * W * H) + (W * H * C // 4 + 8 * W * H // 4)  # read c + write 8 per pixel, both levels
    front_gbs = front_bytes / (kmean["front_ms"] * 1e-3) / 1e9
    traffic = latest_profile_traffic(kname)
    roofline = {
        "kernel": kname, "bound": "l2", "achieved": sim_gbs, "peak": l2_gbs, "unit": "GB/s", "frac": sim_gbs / l2_gbs,
        "traffic": traffic[0] if traffic else None,
        "traffic_source": ("%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one launch, captured at commit %s)" % (traffic[1], traffic[2]))
        if traffic else "no committed capture names this kernel yet",
        "peak_source": "measured in this run: sbm_time_l2_read streams a buffer of the kernel's working-set size (%d bytes) from L2 on every SM "
                       "(ld.global.cg.v4, 592 CTAs, CUDA events).  The kernel's %.1f MB of algorithmic bytes come out of a %.2f MiB L2-resident "
                       "working set (DRAM traffic per launch: see traffic), so the HBM copy peak is not what bounds it" % (lm_bytes, coarse / 1e6, lm_bytes / 2**20),
        "hbm": {"peak": hbm_peak, "frac": sim_gbs / hbm_peak, "peak_source": peak_src},
        "algorithmic_bytes_per_launch": coarse, "avg_launch_ms": sim_ms,
        "bytes_stored_per_evaluation": bits_per_resp / 8.0,
        "how": "50 launches of the pipeline's own coarse-similarity grid, captured as one CUDA graph, between two CUDA events on the "
               "launching stream (sbm_time_similarity); stream launches on this host are CPU-paced at ~5 us (tools/launch_quant.cu)",
        "kernel_ms_warm": kernel_ms,
        "other_kernels": {"front_end(5 launches)": {"ms": kmean["front_ms"], "algorithmic_bytes": front_bytes, "GB/s": front_gbs,
                                                    "frac_of_hbm_peak": front_gbs / hbm_peak},
                          "similarity(+events)": {"ms": kmean["similarity_ms"]},
                          "k_refine": {"ms": kmean["refine_ms"]}, "pipeline_total_ms": kmean["total_ms"],
                          "note": "stage times bracketed by CUDA events inside one frame; each bracket adds a few us"},
    }

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "ms_per_frame": ms_per_step / B, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "step_ms": {"min": min(per_step), "median": float(np.median(per_step)), "max": max(per_step), "note": "rank 0's per-step CUDA-event times"},
        "config": {"workload": M1_NAME, "frames_per_step_per_gpu": B, "sharding": "frames (each rank its own frames, all templates)",
                   "l2_policy": "inputs larger than L2: %d MiB of distinct frames per step per GPU" % (B * W * H * C >> 20),
                   "evals_per_frame": evals_step / (B * world), "final_matches_per_step": n_final // max(1, args.steps),
                   "records_per_frame_max": max_recs, "exchange_bytes_per_rank_per_step": (B * blk) if world > 1 else 0,
                   "exchange": "one ncclAllGather of result blocks per step, issued by sbm_group_match_batch_device" if world > 1 else "none"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * W * H * C, "d2h_bytes_per_step": d2h,
                "ms_per_frame": e2e_s / B * 1e3, "api": ("Detector.match_batch(%d pinned host frames) -> %d final host lists" % (B, B)) if world == 1 else
                ("Detector.group_match_batch(%d pinned host frames per rank) -> the %d final host lists of all ranks, on every rank" % (B, B * world)),
                "host_numa": numa,
                "pcie_floor": {"h2d_us_per_frame": h2d_us_per_frame, "h2d_GBps_per_gpu": W * H * C / (h2d_us_per_frame * 1e-6) / 1e9,
                               "h2d_GBps_all_gpus": world * W * H * C / (h2d_us_per_frame * 1e-6) / 1e9,
                               "e2e_over_floor": (e2e_s / B * 1e6) / h2d_us_per_frame,
                               "how": "the step's pinned frames copied host -> device back to back on one stream, best of 3 (max over ranks); under "
                                      "torchrun every rank does this at the same time, so the floor includes the contention for the host"}},
        "latency": {"device_ms_per_frame_single_stream": serial_ms_per_frame, "match_call_ms_host_to_host": single_call_ms},
        "gpu_launches": launches_per_frame * B * args.steps,
        "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
    }
    if c4:
        line["c4_template_sharded"] = c4
    if c5:
        line["c5_frame_sharded"] = c5
    if _JSON_FD is not None:
        sys.stdout.flush()
        os.write(_JSON_FD, (json.dumps(line) + "\n").encode())
    else:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
