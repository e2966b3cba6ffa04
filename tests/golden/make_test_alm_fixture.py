"""Generates tests/golden/test_alm_8d.json from the reference's own fixture
alpaqa/test/outer/matrices/test-alm.mat.ipp (A, B of ALM.multipleshooting8D) together with the
expected values written in alpaqa/test/outer/test-alm.cpp:291-303.  Run in the build container
(the reference does not travel to the GPU box):   python tests/golden/make_test_alm_fixture.py
"""
import json
import os
import re

REF = "/root/reference/alpaqa/test/outer"
HERE = os.path.dirname(os.path.abspath(__file__))


def parse_matrices(path):
    text = open(path).read()
    out = {}
    for name, body in re.findall(r"(\w)\s*<<\s*([^;]*);", text):
        out[name] = [float(t) for t in re.findall(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?", body)]
    return out


def parse_expected(path):
    text = open(path).read()
    sect = text[text.index("TEST(ALM, multipleshooting8D)"):text.index("TEST(ALM, multipleshooting8D)") + 20000]
    exp = {}
    for name in ("u_ref", "xs_ref", "y_ref"):
        m = re.search(name + r"\s*<<\s*([^;]*);", sect)
        exp[name] = [float(t) for t in re.findall(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?", m.group(1))]
    return exp


if __name__ == "__main__":
    mats = parse_matrices(os.path.join(REF, "matrices", "test-alm.mat.ipp"))
    assert len(mats["A"]) == 64 and len(mats["B"]) == 64
    fix = {"source": "alpaqa/test/outer/matrices/test-alm.mat.ipp + alpaqa/test/outer/test-alm.cpp:181-304",
           "A_rowmajor_8x8": mats["A"], "B_rowmajor_8x8": mats["B"], "tolerance": 6e-4}
    fix.update(parse_expected(os.path.join(REF, "test-alm.cpp")))
    json.dump(fix, open(os.path.join(HERE, "test_alm_8d.json"), "w"), indent=1)
    print("wrote test_alm_8d.json", {k: len(v) for k, v in fix.items() if isinstance(v, list)})
