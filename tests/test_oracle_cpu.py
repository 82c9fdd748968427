"""Pins the oracle (the reference's own CPU code built headless, oracle/_ref/nori_ref) to the
reference's known answers, and the committed fixtures to the oracle."""
import json
import os

import numpy as np
import pytest

from conftest import golden, have_oracle_binary, run_oracle, GOLDEN

needs_oracle = pytest.mark.skipif(not have_oracle_binary(), reason="oracle/_ref/nori_ref not built (make -C oracle ref)")


def test_pcg32_known_answer_file():
    """ext/pcg32/pcg32-demo.out, Round 1 (seed 42, stream 54): 0xa15c02b7 0x7b47f409 ..."""
    line = open(golden("rng_kat.txt")).read().split("\n")[2].split()
    assert line[:6] == ["a15c02b7", "7b47f409", "ba1d3330", "83d2f293", "bfa4784b", "cbed606e"]


@needs_oracle
def test_oracle_rng_matches_fixture():
    assert run_oracle("rng", "16") == open(golden("rng_kat.txt")).read()


@needs_oracle
@pytest.mark.parametrize("name,xml", [("cbox", "scenes/cbox/cbox_path_mis.xml"), ("table", "scenes/table/table_path_mis.xml")])
def test_fixture_hits_reproduce(nb, name, xml, tmp_path):
    """The committed hit goldens are what the oracle computes today on the committed rays."""
    out = tmp_path / "hits.bin"
    run_oracle("hits", golden(xml), golden(name + "_rays.bin"), str(out))
    full = np.fromfile(out, dtype=np.uint8).reshape(-1, 128)
    h = np.fromfile(golden(name + "_hits.bin"), dtype=nb.hit_dtype)
    assert len(full) == len(h)
    assert np.array_equal(full[:, 0:4].copy().view("<i4")[:, 0], h["shape"])
    assert full[:, 8:20].tobytes() == np.stack([h["t"], h["u"], h["v"]], 1).astype("<f4").tobytes()


@needs_oracle
def test_reference_ttest_known_answers():
    """scenes/pa4/tests/test-direct.xml (path_mats x5, path_mis x5): the reference's own
    Student-t references hold for the oracle build (reduced sample count keeps this quick)."""
    ref = "/root/reference/scenes/pa4/tests/test-direct.xml"
    if not os.path.exists(ref):
        pytest.skip("reference tree not mounted")
    lines = [json.loads(l) for l in run_oracle("ttest", ref, "samples=20000").strip().splitlines()]
    assert lines[-1]["passed"] == lines[-1]["total"] == 10
