"""What each kernel of the frame graph costs with several frames in flight: the M1 frame rate (device-resident, 64 frames,
SBM_FRAME_SLOTS slots) with one kernel at a time launched as ONE CTA (SBM_DEBUG_SHRINK, results are garbage then).
    python tools/inflight_shares.py            # runs itself once per kernel"""
import os, subprocess, sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
NAMES = ["orient_L0", "spread_L0", "pyrdown", "orient_L1", "spread_L1", "similarity", "refine"]
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    import torch
    from shape_based_matching_b200 import Detector
    import bench
    base, cid, tdesc, feats, nf = bench.load_workload()
    B = 64
    frames = bench.make_frames(base, B, 1000)
    det = Detector(nf, (4, 8)); det.setClassTemplates(cid, tdesc, feats)
    d = torch.from_numpy(frames).cuda()
    recs = torch.zeros((B, 4096 * 20), dtype=torch.uint8, device="cuda"); cnt = torch.zeros((B, 2), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream()
    def step():
        det.match_batch_device(d.data_ptr(), 1024 * 1024 * 3, B, 1024, 1024, 3, 90.0, [cid], recs.data_ptr(), 4096, cnt.data_ptr(), st.cuda_stream)
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(5): step()
    e1.record(st)
    torch.cuda.synchronize()
    print("%.2f" % (e0.elapsed_time(e1) * 1e3 / (5 * B)))
    sys.exit(0)
def run(mask):
    env = dict(os.environ, SBM_DEBUG_SHRINK=str(mask))
    return float(subprocess.run([sys.executable, __file__, "--one"], env=env, capture_output=True, text=True).stdout.strip().splitlines()[-1])
full = run(0)
print("all kernels: %.2f us per frame in flight" % full)
for i, n in enumerate(NAMES):
    t = run(1 << i)
    print("  without %-11s %.2f  (costs %.2f)" % (n, t, full - t))
t = run(127)
print("  all seven as one CTA each: %.2f (the floor of the graph: node starts, memset, copies, host enqueue)" % t)
