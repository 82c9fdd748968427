"""CPU-side checks of the drop-in boundary: the library loads, exports what include/norib200.h
declares, agrees with the Python mirrors on record sizes, and refuses to run without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden


def test_exports_every_declared_symbol(nb, built_lib):
    hdr = open(os.path.join(ROOT, "include", "norib200.h")).read()
    declared = set(re.findall(r"\b(nb_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(nb.EXPORTS), declared ^ set(nb.EXPORTS)
    for name in declared:
        assert hasattr(built_lib, name), name


def test_record_sizes_match(nb, built_lib):
    assert built_lib.nb_abi_version() == 7
    expect = [nb.shape_dtype.itemsize, nb.bsdf_dtype.itemsize, nb.emitter_dtype.itemsize, nb.camera_dtype.itemsize,
              nb.bvh_node_dtype.itemsize, C.sizeof(nb.SceneDesc), nb.ray_dtype.itemsize, nb.hit_dtype.itemsize,
              nb.render_params_dtype.itemsize, C.sizeof(nb.Counters), nb.medium_dtype.itemsize, nb.texture_dtype.itemsize]
    got = [built_lib.nb_abi_sizeof(i) for i in range(12)]
    assert got == expect
    assert (nb.shape_dtype.itemsize, nb.bsdf_dtype.itemsize, nb.emitter_dtype.itemsize, nb.bvh_node_dtype.itemsize) == (64, 32, 32, 32)
    assert C.sizeof(nb.Camera) == nb.camera_dtype.itemsize == 160


def test_no_cpu_fallback(nb, built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(nb.NbError) as e:
        nb.Context(0)
    assert e.value.code == nb.NB_E_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_missing_library_fails_loudly(nb, tmp_path):
    with pytest.raises(FileNotFoundError):
        nb.load_library(str(tmp_path / "libnorib200.so"))


def test_nbs_roundtrip(nb, tmp_path):
    s = nb.read_nbs(golden("cbox.nbs"))
    assert len(s["shapes"]) == 6 and s["faces"].size == 36 and len(s["bvh_nodes"]) == 3
    assert s["shapes"]["type"].tolist() == [0, 0, 0, 1, 1, 0]
    assert s["bsdfs"]["type"].tolist() == [0, 0, 0, 1, 2, 0]
    p = tmp_path / "x.nbs"
    nb.write_nbs(str(p), s)
    t = nb.read_nbs(str(p))
    assert set(s) == set(t)
    for k in s:
        assert s[k].tobytes() == t[k].tobytes(), k


def test_cli_fails_loudly_without_a_gpu(nb, built_lib):
    """nori_b200 on a machine without a GPU: an error that says so and a non-zero exit code, for scenes and for the
    reference's <test> files alike — no CPU path is taken."""
    import subprocess
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nori_b200")
    golden = os.path.join(ROOT, "tests", "golden", "scenes")
    for args in ([os.path.join(golden, "cbox", "cbox_path_mis_small.xml"), "/tmp/never_written.npy", "spp=1"],
                 [os.path.join(golden, "ttests", "pa3", "ttest-microfacet.xml")]):
        r = subprocess.run([cli] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 1 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr[-300:])
    assert subprocess.run([cli], stdout=subprocess.PIPE, stderr=subprocess.PIPE).returncode == 2
