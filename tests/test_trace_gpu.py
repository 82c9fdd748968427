"""Hit parity (the first gate): the CUDA traversal, called through the C ABI with host buffers,
must return the SAME primitive and bit-identical t,u,v as the reference's BVH::rayIntersect
(src/utils/bvh.cpp:402-460) on fixed ray sets — committed goldens, and (when the oracle binary
travelled with the repo) a larger set generated and answered by the oracle on the spot."""
import numpy as np
import pytest

from conftest import golden, have_oracle_binary, run_oracle

pytestmark = pytest.mark.gpu


def check(nb, ctx, scene, rays, want, flat=-1):
    # 0: the reference's left-first tree walk; 1: flat scan of the primitive list; 2: the ordered (near child
    # first) walk with its tie rule — all three must return the reference's bits
    ctx.set_option("ordered", 1 if flat == 2 else 0)
    flat = 0 if flat == 2 else flat
    ctx.set_option("flat_scan", flat)
    ctx.load(scene)
    got = ctx.trace_rays(rays, any_hit=True)
    assert len(got) == len(want)
    assert np.array_equal(got["shape"], want["shape"]), "shape index mismatch on %d rays" % np.sum(got["shape"] != want["shape"])
    assert np.array_equal(got["prim"], want["prim"]), "primitive index mismatch on %d rays" % np.sum(got["prim"] != want["prim"])
    hit = want["shape"] >= 0
    for k in ("t", "u", "v"):      # bit-exact, not allclose
        assert got[k][hit].tobytes() == want[k][hit].tobytes(), k
    assert np.all(np.isinf(got["t"][~hit]))
    assert np.array_equal(got["any_hit"], want["any_hit"])


@pytest.mark.parametrize("flat", [0, 1, 2])
@pytest.mark.parametrize("name", ["cbox", "table"])
def test_golden_hits_bit_exact(nb, ctx, name, flat):
    scene = nb.read_nbs(golden(name + ".nbs"))
    rays = np.fromfile(golden(name + "_rays.bin"), dtype="<f4").reshape(-1, 8)
    want = np.fromfile(golden(name + "_hits.bin"), dtype=nb.hit_dtype)
    assert len(rays) > 20000
    check(nb, ctx, scene, rays, want, flat)


@pytest.mark.skipif(not have_oracle_binary(), reason="oracle/_ref/nori_ref not present")
@pytest.mark.parametrize("flat", [0, 1, 2])
@pytest.mark.parametrize("name,xml,n", [("cbox", "scenes/cbox/cbox_path_mis.xml", 200000), ("table", "scenes/table/table_path_mis.xml", 150000)])
def test_live_oracle_hits_bit_exact(nb, ctx, name, xml, n, tmp_path, flat):
    """~1M rays per scene: primary, bounce (mint == Epsilon -> adaptive epsilon) and NEE shadow
    rays with finite maxt, answered by the reference code itself on this machine."""
    rays_f, hits_f = str(tmp_path / "rays.bin"), str(tmp_path / "hits.bin")
    run_oracle("genrays", golden(xml), rays_f, str(n), "777")
    run_oracle("hits", golden(xml), rays_f, hits_f)
    rays = np.fromfile(rays_f, dtype="<f4").reshape(-1, 8)
    full = np.fromfile(hits_f, dtype=np.uint8).reshape(-1, 128)
    scene = nb.read_nbs(golden(name + ".nbs"))
    want = np.zeros(len(full), dtype=nb.hit_dtype)
    want["shape"] = full[:, 0:4].copy().view("<i4")[:, 0]
    want["prim"] = full[:, 4:8].copy().view("<i4")[:, 0]
    tuv = full[:, 8:20].copy().view("<f4")
    want["t"], want["u"], want["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
    want["any_hit"] = full[:, 28:32].copy().view("<u4")[:, 0]
    assert len(rays) > 500000
    check(nb, ctx, scene, rays, want, flat)


def test_edge_cases(nb, ctx):
    scene = nb.read_nbs(golden("cbox.nbs"))
    ctx.load(scene)
    assert len(ctx.trace_rays(np.zeros((0, 8), dtype="<f4"))) == 0          # empty batch
    rays = np.array([
        [0, 0.9, 5.4, 0, 0, -1, 1e-4, np.inf],        # axis-aligned: d.x == d.y == 0 branch of the slab test
        [0, 0.9, 5.4, 0, 0, -1, 1e-4, 1e-3],          # maxt before anything
        [0, 0.9, 5.4, 0, 0, -1, 5.0, 1.0],            # maxt < mint: rejected up front (bvh.cpp:412)
        [0, 0.9, 5.4, 0, 0, 1, 1e-4, np.inf],         # looking away: miss
        [0, 50, 0, 0, -1, 0, 1e-4, np.inf],           # from above through the ceiling light
    ], dtype="<f4")
    h = ctx.trace_rays(rays, any_hit=True)
    assert h["shape"][0] >= 0 and h["any_hit"][0] == 1
    assert h["shape"][1] == -1 and h["shape"][2] == -1 and h["shape"][3] == -1
    assert h["any_hit"][1] == 0 and h["any_hit"][2] == 0 and h["any_hit"][3] == 0
    assert h["shape"][4] >= 0


def test_errors(nb, built_lib):
    c = nb.Context(0)
    try:
        with pytest.raises(nb.NbError) as e:
            c.trace_rays(np.zeros((1, 8), dtype="<f4"))
        assert e.value.code == nb.NB_E_STATE
        scene = nb.read_nbs(golden("cbox.nbs"))
        bad = dict(scene); bad["bsdfs"] = scene["bsdfs"].copy(); bad["bsdfs"]["type"][0] = 77
        with pytest.raises(nb.NbError) as e:
            c.upload_scene(bad)
        assert e.value.code == nb.NB_E_UNSUPPORTED      # caller falls back to its CPU renderer
        c.upload_scene(scene)
        with pytest.raises(nb.NbError) as e:
            c.build_bvh(scene["bvh_nodes"], scene["bvh_indices"][:-1])
        assert e.value.code == nb.NB_E_INVALID
    finally:
        c.close()
