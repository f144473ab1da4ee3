"""A C++ caller sharding Detector::match over GPUs (tests/cpp/group_driver.cpp: Detector::setGroup, no Python / torch / MPI
in the ranks).  With one GPU a group of one rank runs the whole path; with two (gpurun --gpus 2) two processes shard case1's
361 templates, or their frames, for real.  Expected: the reference's golden list."""
import os
import subprocess

import numpy as np
import pytest

from _fixtures import load_case, load_templates, write_templates_yaml
from test_cpp_dropin import golden_tuples, parse_matches

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tests", "cpp", "group_driver")


@pytest.mark.parametrize("n_ranks", [1, 2])
@pytest.mark.parametrize("mode", [0, 1, 2], ids=["templates", "templates_contiguous", "frames"])
def test_cpp_group_equals_reference(n_ranks, mode, tmp_path):
    import torch

    if torch.cuda.device_count() < n_ranks:
        pytest.skip("needs %d GPUs (gpurun --gpus %d)" % (n_ranks, n_ranks))
    assert os.path.exists(DRIVER), "build it: python -m shape_based_matching_b200.build"
    case = load_case("case1")
    cid, tdesc, feats, nf = load_templates(case["templates"])
    write_templates_yaml(str(tmp_path / ("%s.yaml" % cid)), cid, tdesc, feats)
    img = np.ascontiguousarray(case["img"])
    img.tofile(tmp_path / "img.bin")
    h, w, c = img.shape
    tail = [str(tmp_path / "%s.yaml"), cid, str(tmp_path / "img.bin"), w, h, c, float(case["threshold"]), nf] + [int(t) for t in case["T"]]
    procs = [subprocess.Popen([DRIVER, str(r), str(n_ranks), str(mode), str(tmp_path / "nccl_id.bin")] + [str(a) for a in tail],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(n_ranks)]
    want = golden_tuples(case["matches"], cid)
    for r, p in enumerate(procs):
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0, out[-3000:]
        assert ("rank %d of %d templates %d" % (r, n_ranks, tdesc.shape[0])) in out  # the caller still sees the full class
        lists = parse_matches(out)
        if mode != 2:
            assert len(lists) == 4
            for got in lists:
                assert got == want
        else:
            assert len(lists) == n_ranks
            for q in range(0, n_ranks, 2):  # even ranks matched the unmirrored frame
                assert lists[q] == want
