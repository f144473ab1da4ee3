"""Device times of BASELINE.json configs 4 and 5 (parity for them is in tests/test_gpu_large.py; bench.py's line is M1).

    python tools/config_times.py            # prints one JSON line per configuration

C5: 2048x2048x1 frames x 720 templates (128/64 features).  C4: 4096x4096x1 frame x 3600 templates (8191/4095
features).  Templates are cut from the frame's own quantised orientations (GPU front end), so every template
has a true match.  Times: CUDA events around sbm_match_device (frame resident in HBM, results left in HBM), and
the coarse similarity kernel alone (sbm_time_similarity)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch

    from test_gpu_large import cut_templates, dense_image

    import shape_based_matching_b200._capi as capi
    if len(sys.argv) > 1:
        capi.LIB_PATH = sys.argv[1]  # a variant build of the library
    from shape_based_matching_b200 import DEVICE_REC, Detector

    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    out = []
    for name, hw, n_t, nf, win, shapes, n_org, thr in (("C5", 2048, 720, (128, 64), 256, 700, 24, 90.0),
                                                       ("C4", 4096, 3600, (8191, 4095), 768, 5000, 36, 80.0)):
        rng = np.random.default_rng(hw)
        img = dense_image(rng, hw, hw, 1, shapes)
        det = Detector(nf[0], (4, 8), 30.0, 60.0)
        quant = [det.quantized_orientations(img, l)[2] for l in range(2)]
        hi = (hw - win - 32) // 16
        origins = [(int(x) * 16, int(y) * 16) for x, y in rng.integers(2, hi, (n_org, 2))]
        tdesc, feats = cut_templates(rng, quant, n_t, win, nf, origins)
        det.setClassTemplates("c", tdesc, feats)
        d_img = torch.from_numpy(img).cuda()
        cap = 1 << 16
        recs = torch.zeros(cap * DEVICE_REC.itemsize, dtype=torch.uint8, device="cuda")
        cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
        stream = torch.cuda.Stream()
        torch.cuda.set_stream(stream)

        def run():
            det.match_device(d_img.data_ptr(), hw, hw, 1, thr, ["c"], d_records=recs.data_ptr(), cap=cap, d_count=cnt.data_ptr(),
                             stream=stream.cuda_stream)

        det.match_records(img, thr, ["c"], cap=cap)  # sizes every buffer, gives the evaluation counts
        st = det.last_stats()
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record(stream)
        for _ in range(iters):
            run()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        sim_ms = det.time_similarity(thr, iters=5)
        evals = st["coarse_evals"] + st["local_evals"]
        t0 = time.perf_counter()
        n = len(det.match_records(img, thr, ["c"], cap=cap))
        call_ms = (time.perf_counter() - t0) * 1e3
        line = {"config": name, "frame": "%dx%dx1" % (hw, hw), "templates": n_t, "features": list(nf), "coarse_evals": st["coarse_evals"],
                "local_evals": st["local_evals"], "records": n, "device_ms_per_frame": ms, "evals_per_s": evals / (ms * 1e-3),
                "k_similarity_ms": sim_ms, "k_similarity_GBps": st["coarse_evals"] / (sim_ms * 1e-3) / 1e9,
                "k_similarity_frac_of_hbm_peak": st["coarse_evals"] / (sim_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "coarse_linear_memory_bytes": 8 * (hw // 2) * (hw // 2), "match_call_ms_host_to_host": call_ms}
        print(json.dumps(line), flush=True)
        out.append(line)
        det.close()
    return out


if __name__ == "__main__":
    main()
