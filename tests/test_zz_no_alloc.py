"""nb_trace_rays_static — the trace entry that allocates nothing (fixed __device__ scratch inside the library,
include/norib200.h) — against the existing nb_trace_rays, through the C ABI, bit for bit, on calls in a row that
share the scratch. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def same(nb, got, want):
    assert got.dtype == nb.hit_dtype and len(got) == len(want)
    assert got.tobytes() == want.tobytes()          # every field of every record, bit for bit (integer + exact fp32 work)


@pytest.mark.parametrize("name", ["cbox", "table"])       # flat scan (14 primitives) and the tree walk (22 766 triangles)
def test_static_entry_matches_trace_rays_on_two_calls_in_a_row(nb, ctx, name):
    scene = nb.read_nbs(golden(name + ".nbs"))
    rays = np.fromfile(golden(name + "_rays.bin"), dtype="<f4").reshape(-1, 8)
    assert nb.TRACE_STATIC_MAX_RAYS >= len(rays) > 20000
    ctx.load(scene)
    want_all = ctx.trace_rays(rays, any_hit=True)
    want_golden = np.fromfile(golden(name + "_hits.bin"), dtype=nb.hit_dtype)
    hit = want_golden["shape"] >= 0
    assert np.array_equal(want_all["shape"], want_golden["shape"]) and want_all["t"][hit].tobytes() == want_golden["t"][hit].tobytes()
    # call 1: the whole set, closest + shadow query
    same(nb, ctx.trace_rays_static(rays, any_hit=True), want_all)
    # call 2, straight after, on the same scratch: a shorter, reordered batch without the shadow query — nothing of
    # call 1 (its longer tail, its any_hit words) may show through
    sub = np.ascontiguousarray(rays[4999::-1])
    want_sub = ctx.trace_rays(sub, any_hit=False)
    got_sub = ctx.trace_rays_static(sub, any_hit=False)
    same(nb, got_sub, want_sub)
    assert not got_sub["any_hit"].any()
    # call 3: the first batch again
    same(nb, ctx.trace_rays_static(rays, any_hit=True), want_all)


def test_static_entry_ordered_walk_and_edges(nb, ctx):
    scene = nb.read_nbs(golden("table.nbs"))
    rays = np.fromfile(golden("table_rays.bin"), dtype="<f4").reshape(-1, 8)[:8192]
    ctx.set_option("ordered", 1)
    ctx.load(scene)
    same(nb, ctx.trace_rays_static(rays, any_hit=True), ctx.trace_rays(rays, any_hit=True))
    assert len(ctx.trace_rays_static(np.zeros((0, 8), dtype="<f4"))) == 0       # empty batch


def test_static_entry_rejects_more_than_its_scratch_holds(nb, ctx):
    ctx.load(nb.read_nbs(golden("cbox.nbs")))
    n = nb.TRACE_STATIC_MAX_RAYS + 1
    rays = np.zeros((n, 8), dtype="<f4")
    out = np.zeros(n, dtype=nb.hit_dtype)
    rc = ctx.lib.nb_trace_rays_static(ctx.handle, rays.ctypes.data_as(C.c_void_p), n, 0, out.ctypes.data_as(C.c_void_p))
    assert rc == nb.NB_E_INVALID
    assert b"NB_TRACE_STATIC_MAX_RAYS" in ctx.lib.nb_last_error(ctx.handle)
    # exactly the maximum is accepted
    full = np.tile(np.fromfile(golden("cbox_rays.bin"), dtype="<f4").reshape(-1, 8), (3, 1))[:nb.TRACE_STATIC_MAX_RAYS]
    assert len(full) == nb.TRACE_STATIC_MAX_RAYS
    same(nb, ctx.trace_rays_static(full, any_hit=True), ctx.trace_rays(full, any_hit=True))
