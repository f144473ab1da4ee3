"""The drop-in C++ API (include/line2Dup.h + libline2Dup_b200.so) driven from C++ (tests/cpp/dropin_driver.cpp):
persistence on the CPU; match / matchBatch / addTemplate / addTemplate_rotate on the GPU against the reference."""
import os
import subprocess

import numpy as np
import pytest

from _fixtures import load_case, load_templates, parse_driver_templates, write_templates_yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tests", "cpp", "dropin_driver")


def run(*args):
    assert os.path.exists(DRIVER), "build it: python -m shape_based_matching_b200.build"
    r = subprocess.run([DRIVER] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    return r.stdout


def parse_matches(text):
    lists, cur = [], None
    for line in text.splitlines():
        if line.startswith("matches "):
            cur = []
            lists.append(cur)
        elif line.startswith("M "):
            p = line.split()
            cur.append((int(p[1]), int(p[2]), int(p[3], 16), p[4], int(p[5])))
    return lists


def golden_tuples(recs, cid):
    return [(int(r["x"]), int(r["y"]), int(np.float32(r["similarity"]).view(np.uint32)), cid, int(r["template_id"])) for r in recs]


@pytest.mark.parametrize("name", ["case0", "case2"])
def test_yaml_read_write_roundtrip(name, tmp_path):
    cid, tdesc, feats, _nf = load_templates(name)
    src = tmp_path / ("%s_in.yaml" % cid)
    write_templates_yaml(str(src), cid, tdesc, feats)
    out = run("yaml", str(tmp_path / "%s_in.yaml"), cid, str(tmp_path / "%s_out.yaml.gz"))
    td, ff = parse_driver_templates(out)
    assert np.array_equal(td, tdesc) and np.array_equal(ff, feats)
    assert "reread %d" % tdesc.shape[0] in out
    assert os.path.getsize(tmp_path / ("%s_out.yaml.gz" % cid)) > 100


@pytest.mark.skipif(not os.path.exists("/root/reference/test/case1/test_templ.yaml"), reason="reference tree not present")
def test_reads_the_reference_fixture_files(tmp_path):
    for case, fmt, cid in (("case1", "/root/reference/test/case1/%s_templ.yaml", "test"),
                           ("case0", "/root/reference/test/case0/%s_templ.yaml", "circle")):
        gcid, tdesc, feats, _nf = load_templates(case)
        assert gcid == cid
        td, ff = parse_driver_templates(run("yaml", fmt, cid, str(tmp_path / "%s.yaml")))
        assert np.array_equal(td, tdesc) and np.array_equal(ff, feats)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["case0_3", "case1"])
def test_cpp_match_equals_reference(name, tmp_path):
    case = load_case(name)
    cid, tdesc, feats, nf = load_templates(case["templates"])
    write_templates_yaml(str(tmp_path / ("%s.yaml" % cid)), cid, tdesc, feats)
    img = np.ascontiguousarray(case["img"])
    img.tofile(tmp_path / "img.bin")
    h, w, c = img.shape
    args = [str(tmp_path / "%s.yaml"), cid, str(tmp_path / "img.bin"), w, h, c, float(case["threshold"]), nf] + [int(t) for t in case["T"]]
    want = golden_tuples(case["matches"], cid)
    lists = parse_matches(run("match", *args))
    assert lists[0] == want  # the requested class
    assert lists[1] == want  # no class list = all classes
    assert lists[2] == []  # unknown class ids are skipped silently (line2Dup.cpp:1136-1138)
    for got in parse_matches(run("batch", *args)):
        assert got == want


@pytest.mark.gpu
def test_cpp_training_equals_reference(tmp_path, ref_available):
    if not ref_available:
        pytest.skip("oracle/_ref not built")
    from oracle import ref

    case = load_case("case1")
    img = np.ascontiguousarray(case["img"][300:680, 310:700])  # the object of the scene and some background
    h, w, c = img.shape
    img.tofile(tmp_path / "train.bin")
    mask = np.zeros((h, w), np.uint8)
    mask[10:-10, 12:-8] = 255
    mask.tofile(tmp_path / "mask.bin")
    for use_mask in (False, True):
        rd = ref.RefDetector(128, (4, 8))
        tid = rd.add_template(img, "trained", mask if use_mask else None)
        assert tid == 0
        for k in range(1, 4):
            assert rd.add_template_rotate("trained", 0, 7.5 * k, w / 2.0, h / 2.0) == k
        want_td, want_ff = rd.export_class("trained")
        out = run("train", str(tmp_path / "train.bin"), w, h, c, str(tmp_path / "mask.bin") if use_mask else "-", 128,
                  str(tmp_path / "%s_trained.yaml"), 3, 7.5, 4, 8)
        assert "addTemplate 0" in out
        td, ff = parse_driver_templates(out)
        assert np.array_equal(td, want_td), (td[:1], want_td[:1])
        assert np.array_equal(ff, want_ff)
