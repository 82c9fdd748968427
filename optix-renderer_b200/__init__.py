"""nori-b200: ctypes binding of the C ABI in include/norib200.h (libnorib200.so).

The library is the product; this module only loads it, mirrors its records and moves numpy
buffers across the boundary. There is no Python or CPU implementation of the path behind it: if
the shared library is missing, or no CUDA device is present, calls fail loudly.

Import with ``importlib.import_module("optix-renderer_b200")`` (the directory name carries a
hyphen) or through ``norib200.py`` at the repo root.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libnorib200.so")
HOST_LIB_PATH = os.path.join(HERE, "libnorib200_host.so")

NB_OK, NB_E_INVALID, NB_E_CUDA, NB_E_NOMEM, NB_E_STATE, NB_E_UNSUPPORTED, NB_E_NO_DEVICE = 0, -1, -2, -3, -4, -5, -6
SHAPE_MESH, SHAPE_SPHERE = 0, 1
BSDF_DIFFUSE, BSDF_MIRROR, BSDF_DIELECTRIC, BSDF_MICROFACET, BSDF_DISNEY = 0, 1, 2, 3, 4
EMITTER_AREA, EMITTER_POINT, EMITTER_ENVMAP, EMITTER_SPOT, EMITTER_DIRECTIONAL, EMITTER_VOLUME = 0, 1, 2, 3, 4, 5
INTEGRATORS = {"path_mis": 0, "path_mats": 1, "direct_ems": 2, "direct_mats": 3, "direct_mis": 4,
               "normals": 5, "direct": 6, "av": 7, "path_vol_mats": 8, "path_vol_mis": 9, "preview": 10}
MEDIUM_VACUUM, MEDIUM_HOMOGENEOUS, MEDIUM_HETEROGENEOUS, MEDIUM_HETEROGENEOUS_BRICKS = 0, 1, 2, 3
PHASE_ISOTROPIC, PHASE_HENYEY_GREENSTEIN, PHASE_SCHLICK = 0, 1, 2
FILTER_RESOLUTION = 32
TRACE_STATIC_MAX_RAYS = 65536
COMM_ID_BYTES = 128

# numpy mirrors of the records (field order and padding as in the header)
shape_dtype = np.dtype([("type", "<u4"), ("prim_count", "<u4"), ("vertex_offset", "<u4"), ("vertex_count", "<u4"),
                        ("face_offset", "<u4"), ("flags", "<u4"), ("bsdf", "<i4"), ("emitter", "<i4"),
                        ("medium", "<i4"), ("cdf_offset", "<u4"), ("pdf_norm", "<f4"), ("radius", "<f4"),
                        ("center", "<f4", 3), ("normal_map", "<u4")])
bsdf_dtype = np.dtype([("type", "<u4"), ("albedo", "<f4", 3), ("alpha", "<f4"), ("int_ior", "<f4"),
                       ("ext_ior", "<f4"), ("ks", "<f4")])
emitter_dtype = np.dtype([("type", "<u4"), ("shape", "<i4"), ("radiance", "<f4", 3), ("position", "<f4", 3)])
camera_dtype = np.dtype([("sample_to_camera", "<f4", 16), ("camera_to_world", "<f4", 16),
                         ("inv_output_size", "<f4", 2), ("near_clip", "<f4"), ("far_clip", "<f4"),
                         ("lens_radius", "<f4"), ("focal_distance", "<f4"), ("width", "<u4"), ("height", "<u4")])
medium_dtype = np.dtype([("type", "<u4"), ("mu_t", "<f4", 3), ("phase", "<u4"), ("phase_param", "<f4"),
                         ("density_scale", "<f4"), ("max_density", "<f4"), ("grid_dims", "<u4", 3),
                         ("first_voxel", "<u4"), ("world_to_grid", "<f4", 12)])
texture_dtype = np.dtype([("type", "<u4"), ("flags", "<u4"), ("width", "<u4"), ("height", "<u4"), ("first_texel", "<u4"),
                          ("value", "<f4", 3), ("value2", "<f4", 3), ("scale", "<f4", 2), ("offset", "<f4", 2),
                          ("rotation", "<f4", 9), ("intensity", "<f4"), ("pad", "<u4")])
bvh_node_dtype = np.dtype([("word0", "<u4"), ("word1", "<u4"), ("bmin", "<f4", 3), ("bmax", "<f4", 3)])
ray_dtype = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("mint", "<f4"), ("maxt", "<f4")])
hit_dtype = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("shape", "<i4"), ("prim", "<i4"), ("any_hit", "<u4")])
render_params_dtype = np.dtype([("integrator", "<u4"), ("filter_radius", "<f4"),
                                ("filter_table", "<f4", FILTER_RESOLUTION + 1), ("pad0", "<u4"), ("seed", "<u8"),
                                ("pool_paths", "<u4"), ("max_depth", "<u4"), ("integrator_params", "<f4", 3),
                                ("layers", "<u4")])
SECTION_DTYPES = {"shapes": shape_dtype, "bsdfs": bsdf_dtype, "emitters": emitter_dtype, "camera": camera_dtype,
                  "bvh_nodes": bvh_node_dtype, "render": render_params_dtype, "media": medium_dtype, "textures": texture_dtype}


class Camera(C.Structure):
    _fields_ = [("sample_to_camera", C.c_float * 16), ("camera_to_world", C.c_float * 16),
                ("inv_output_size", C.c_float * 2), ("near_clip", C.c_float), ("far_clip", C.c_float),
                ("lens_radius", C.c_float), ("focal_distance", C.c_float), ("width", C.c_uint32), ("height", C.c_uint32)]


class SceneDesc(C.Structure):
    _fields_ = [("n_shapes", C.c_uint32), ("shapes", C.c_void_p),
                ("n_vertices", C.c_uint64), ("positions", C.c_void_p), ("normals", C.c_void_p), ("uvs", C.c_void_p),
                ("tangents", C.c_void_p), ("bitangents", C.c_void_p),
                ("n_faces", C.c_uint64), ("faces", C.c_void_p),
                ("n_bsdfs", C.c_uint32), ("bsdfs", C.c_void_p),
                ("n_emitters", C.c_uint32), ("emitters", C.c_void_p), ("emitter_cdf", C.c_void_p),
                ("n_area_cdf", C.c_uint64), ("area_cdf", C.c_void_p),
                ("camera", Camera),
                ("n_media", C.c_uint32), ("media", C.c_void_p), ("ambient_medium", C.c_int32),
                ("n_density", C.c_uint64), ("density", C.c_void_p),
                ("n_textures", C.c_uint32), ("textures", C.c_void_p),
                ("n_texels", C.c_uint64), ("texels", C.c_void_p),
                ("n_params", C.c_uint64), ("params", C.c_void_p)]


class BvhImport(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("nodes", C.c_void_p), ("n_indices", C.c_uint32), ("indices", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays_closest", C.c_uint64), ("rays_any", C.c_uint64),
                ("iterations", C.c_uint64), ("kernel_launches", C.c_uint64), ("invalid_samples", C.c_uint64),
                ("device_ms", C.c_double), ("trace_ms", C.c_double), ("shade_ms", C.c_double),
                ("regen_ms", C.c_double)]


EXPORTS = ["nb_abi_version", "nb_abi_sizeof", "nb_create", "nb_destroy", "nb_last_error", "nb_upload_scene",
           "nb_build_bvh", "nb_trace_rays", "nb_trace_rays_device", "nb_render_begin", "nb_render_subframe", "nb_render_adaptive_round", "nb_adaptive_block_passes",
           "nb_sync", "nb_film_clear", "nb_film_shape", "nb_film_download", "nb_film_device_ptr",
           "nb_get_counters", "nb_set_option", "nb_sample_slice", "nb_bvh_stats", "nb_medium_probe", "nb_emitter_probe",
           "nb_bsdf_probe", "nb_film_download_layer", "nb_trace_rays_static", "nb_create_multi", "nb_comm_unique_id",
           "nb_comm_init_rank", "nb_comm_info", "nb_film_reduce", "nb_timer_mark", "nb_timer_elapsed", "nb_rng_probe"]

_lib = None


class NbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("norib200 error %d: %s" % (code, text))
        self.code = code


def load_library(path=None):
    """dlopen libnorib200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("NORIB200_LIB") or LIB_PATH      # NORIB200_LIB: a tuning build of the same library
    if not os.path.exists(path):
        raise FileNotFoundError("%s is not built: run `python optix-renderer_b200/build.py` (there is no "
                                "CPU or Python fallback for the render path)" % path)
    lib = C.CDLL(path)
    lib.nb_last_error.restype = C.c_char_p
    lib.nb_last_error.argtypes = [C.c_void_p]
    lib.nb_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.nb_destroy.argtypes = [C.c_void_p]
    lib.nb_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
    lib.nb_comm_unique_id.argtypes = [C.c_void_p]
    lib.nb_comm_init_rank.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.nb_comm_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.nb_film_reduce.argtypes = [C.c_void_p]
    lib.nb_timer_mark.argtypes = [C.c_void_p, C.c_int]
    lib.nb_rng_probe.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
    lib.nb_timer_elapsed.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.nb_upload_scene.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
    lib.nb_build_bvh.argtypes = [C.c_void_p, C.c_int, C.POINTER(BvhImport)]
    lib.nb_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    lib.nb_trace_rays_static.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    lib.nb_trace_rays_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_float)]
    lib.nb_render_begin.argtypes = [C.c_void_p, C.c_void_p]
    lib.nb_render_subframe.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    lib.nb_render_adaptive_round.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
    lib.nb_adaptive_block_passes.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    lib.nb_sync.argtypes = [C.c_void_p]
    lib.nb_film_clear.argtypes = [C.c_void_p]
    lib.nb_film_shape.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.nb_film_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    lib.nb_film_download_layer.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.nb_film_device_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.nb_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
    lib.nb_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    lib.nb_sample_slice.argtypes = [C.c_uint32] * 4 + [C.POINTER(C.c_uint32)] * 2
    lib.nb_bvh_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_double)]
    lib.nb_emitter_probe.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.nb_bsdf_probe.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.nb_medium_probe.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64, C.c_void_p]
    if path in (LIB_PATH, os.environ.get("NORIB200_LIB")):
        _lib = lib
    return lib


# ---- .nbs container (include/nb_scene_file.h) ---------------------------------------------------
NBS_MAGIC = 0x3153424E
_sec_dtype = np.dtype([("name", "S16"), ("elem_size", "<u4"), ("reserved", "<u4"), ("count", "<u8"), ("offset", "<u8")])


def read_nbs(path):
    """Returns {section name: numpy array}. Record sections come back as structured arrays."""
    raw = np.fromfile(path, dtype=np.uint8)
    magic, n = np.frombuffer(raw[:8].tobytes(), dtype="<u4")
    if magic != NBS_MAGIC:
        raise ValueError("%s: not an .nbs file" % path)
    table = np.frombuffer(raw[8:8 + 40 * int(n)].tobytes(), dtype=_sec_dtype)
    out = {}
    for sec in table:
        name = sec["name"].decode().rstrip("\0")
        a, b = int(sec["offset"]), int(sec["offset"]) + int(sec["count"]) * int(sec["elem_size"])
        buf = raw[a:b].tobytes()
        if name in SECTION_DTYPES:
            dt = SECTION_DTYPES[name]
            assert dt.itemsize == int(sec["elem_size"]), (name, dt.itemsize, int(sec["elem_size"]))
            out[name] = np.frombuffer(buf, dtype=dt).copy()
        elif name in ("faces", "bvh_indices", "spp", "adaptive"):
            out[name] = np.frombuffer(buf, dtype="<u4").copy()
        elif name == "ambient":
            out[name] = np.frombuffer(buf, dtype="<i4").copy()
        else:
            out[name] = np.frombuffer(buf, dtype="<f4").copy()
    return out


def write_nbs(path, sections):
    names = list(sections)
    arrays = [np.ascontiguousarray(sections[k]) for k in names]
    table = np.zeros(len(names), dtype=_sec_dtype)
    off = 8 + 40 * len(names)
    for i, (k, a) in enumerate(zip(names, arrays)):
        off = (off + 15) & ~15
        table[i] = (k.encode(), a.dtype.itemsize, 0, a.size, off)
        off += a.nbytes
    with open(path, "wb") as f:
        f.write(np.array([NBS_MAGIC, len(names)], dtype="<u4").tobytes())
        f.write(table.tobytes())
        for i, a in enumerate(arrays):
            f.write(b"\0" * (int(table[i]["offset"]) - f.tell()))
            f.write(a.tobytes())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class Context:
    """One nb_ctx: a device, its uploaded scene and (optionally) a render in progress."""

    def __init__(self, device=0, _handle=None):
        self.lib = load_library()
        self.handle = C.c_void_p()
        if _handle is not None:                 # one member of a create_multi() group
            self.handle = C.c_void_p(_handle)
        else:
            rc = self.lib.nb_create(device, C.byref(self.handle))
            if rc != NB_OK:
                raise NbError(rc, self.lib.nb_last_error(None).decode())
        self.device = device
        self._keep = []

    def comm_init(self, n_ranks, rank, unique_id):
        """nb_comm_init_rank: join the NCCL communicator of a one-process-per-GPU job (collective)."""
        buf = C.create_string_buffer(bytes(unique_id), COMM_ID_BYTES)
        self._check(self.lib.nb_comm_init_rank(self.handle, n_ranks, rank, buf))

    def comm_info(self):
        r, n, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.nb_comm_info(self.handle, C.byref(r), C.byref(n), C.byref(v)))
        return {"rank": r.value, "n_ranks": n.value, "nccl_version": v.value}

    def timer_mark(self, which):
        self._check(self.lib.nb_timer_mark(self.handle, which))

    def timer_elapsed(self):
        """CUDA-event milliseconds between timer_mark(0) and timer_mark(1) on the library's stream (waits for mark 1)."""
        ms = C.c_float()
        self._check(self.lib.nb_timer_elapsed(self.handle, C.byref(ms)))
        return ms.value

    def film_reduce(self):
        """nb_film_reduce: ncclReduce(sum) of the film accumulators to rank 0, queued on the context's stream."""
        self._check(self.lib.nb_film_reduce(self.handle))

    def _check(self, rc):
        if rc != NB_OK:
            raise NbError(rc, self.lib.nb_last_error(self.handle).decode())

    def close(self):
        if self.handle:
            self.lib.nb_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload_scene(self, s):
        """s: dict of sections as read_nbs returns them (bvh_* used by build_bvh)."""
        d = SceneDesc()
        keep = {k: np.ascontiguousarray(v) for k, v in s.items()}
        d.n_shapes = len(keep["shapes"]); d.shapes = _ptr(keep["shapes"])
        d.n_vertices = keep["positions"].size // 3; d.positions = _ptr(keep["positions"])
        d.normals = _ptr(keep.get("normals")); d.uvs = _ptr(keep.get("uvs"))
        d.tangents = _ptr(keep.get("tangents")); d.bitangents = _ptr(keep.get("bitangents"))
        d.n_faces = keep["faces"].size // 3; d.faces = _ptr(keep["faces"])
        bs = keep.get("bsdfs"); d.n_bsdfs = 0 if bs is None else len(bs); d.bsdfs = _ptr(bs)
        em = keep.get("emitters"); d.n_emitters = 0 if em is None else len(em); d.emitters = _ptr(em)
        d.emitter_cdf = _ptr(keep.get("emitter_cdf")) if d.n_emitters else None
        ac = keep.get("area_cdf"); d.n_area_cdf = 0 if ac is None else ac.size; d.area_cdf = _ptr(ac)
        if "camera" in keep:
            C.memmove(C.byref(d.camera), keep["camera"].ctypes.data, C.sizeof(Camera))
        md = keep.get("media"); d.n_media = 0 if md is None else len(md); d.media = _ptr(md)
        d.ambient_medium = int(keep["ambient"][0]) if "ambient" in keep else -1
        dn = keep.get("density"); d.n_density = 0 if dn is None else dn.size; d.density = _ptr(dn)
        tx = keep.get("textures"); d.n_textures = 0 if tx is None else len(tx); d.textures = _ptr(tx)
        tl = keep.get("texels"); d.n_texels = 0 if tl is None else tl.size // 4; d.texels = _ptr(tl)
        pr = keep.get("params"); d.n_params = 0 if pr is None else pr.size; d.params = _ptr(pr)
        self._check(self.lib.nb_upload_scene(self.handle, C.byref(d)))

    def build_bvh(self, nodes, indices):
        nodes = np.ascontiguousarray(nodes); indices = np.ascontiguousarray(indices, dtype="<u4")
        imp = BvhImport(len(nodes), _ptr(nodes), len(indices), _ptr(indices))
        self._check(self.lib.nb_build_bvh(self.handle, 0, C.byref(imp)))

    def build_bvh_device(self, sah=False):
        """K_bvh_build: the tree built on the GPU from the uploaded scene (no host tree needed) — a linear BVH (2 ms per
        million triangles) or, sah=True, the reference's SAH criterion top-down (NB_BVH_DEVICE_SAH)."""
        self._check(self.lib.nb_build_bvh(self.handle, 2 if sah else 1, None))

    def bvh_stats(self):
        n, m, ms = C.c_uint32(), C.c_uint32(), C.c_double()
        self._check(self.lib.nb_bvh_stats(self.handle, C.byref(n), C.byref(m), C.byref(ms)))
        return {"nodes": n.value, "slots": m.value, "build_ms": ms.value}

    def load(self, s, device_bvh=False):
        self.upload_scene(s)
        if device_bvh:
            self.build_bvh_device(sah=(device_bvh == "sah"))
        else:
            self.build_bvh(s["bvh_nodes"], s["bvh_indices"])

    def emitter_probe(self, emitter, records):
        """records: (n, 5) float32 [ref.xyz, sample.xy]; returns (n, 12) [wi, value, pdf, maxt, eval, 0]."""
        rec = np.ascontiguousarray(records, dtype="<f4").reshape(-1, 5)
        out = np.zeros((len(rec), 12), dtype="<f4")
        self._check(self.lib.nb_emitter_probe(self.handle, emitter, _ptr(rec), len(rec), _ptr(out)))
        return out

    def bsdf_probe(self, bsdf, records):
        """records: (n, 10) float32 [wi.xyz, wo.xyz, sample.xy, uv.xy]; returns (n, 12)
        [eval rgb, pdf, sample weight rgb, sampled wo xyz, discrete, pdf of the sampled direction]."""
        rec = np.ascontiguousarray(records, dtype="<f4").reshape(-1, 10)
        out = np.zeros((len(rec), 12), dtype="<f4")
        self._check(self.lib.nb_bsdf_probe(self.handle, bsdf, _ptr(rec), len(rec), _ptr(out)))
        return out

    def rng_probe(self, initstate, initseq, n):
        """(n nextUInt(), n nextFloat()) of the device pcg32 seeded with seed(initstate, initseq)."""
        u, f = np.zeros(n, dtype="<u4"), np.zeros(n, dtype="<f4")
        self._check(self.lib.nb_rng_probe(self.handle, initstate, initseq, n, _ptr(u), _ptr(f)))
        return u, f

    def medium_probe(self, medium, rays, samples=1024, seed=1):
        """rays: (n, 8) float32 [o, d, mint, maxt]; returns (n, 4) [density(o), P(free path < maxt), mean path, mean Tr]."""
        rays = np.ascontiguousarray(rays, dtype="<f4").reshape(-1, 8)
        out = np.zeros((len(rays), 4), dtype="<f4")
        self._check(self.lib.nb_medium_probe(self.handle, medium, _ptr(rays), len(rays), samples, seed, _ptr(out)))
        return out

    def trace_rays(self, rays, any_hit=False):
        """rays: (n, 8) float32 [o, d, mint, maxt] or ray_dtype; returns hit_dtype array."""
        rays = np.ascontiguousarray(rays)
        n = rays.size * rays.dtype.itemsize // 32
        out = np.zeros(n, dtype=hit_dtype)
        self._check(self.lib.nb_trace_rays(self.handle, _ptr(rays), n, int(any_hit), _ptr(out)))
        return out

    def trace_rays_static(self, rays, any_hit=False):
        """nb_trace_rays_static: the same query through the library's fixed device scratch (no allocation,
        at most TRACE_STATIC_MAX_RAYS rays per call)."""
        rays = np.ascontiguousarray(rays)
        n = rays.size * rays.dtype.itemsize // 32
        out = np.zeros(n, dtype=hit_dtype)
        self._check(self.lib.nb_trace_rays_static(self.handle, _ptr(rays), n, int(any_hit), _ptr(out)))
        return out

    def trace_rays_device(self, rays_ptr, n, out_ptr, any_hit=False, repeats=1):
        ms = C.c_float()
        self._check(self.lib.nb_trace_rays_device(self.handle, C.c_void_p(rays_ptr), n, int(any_hit),
                                                  C.c_void_p(out_ptr), repeats, C.byref(ms)))
        return ms.value

    def render_begin(self, params):
        params = np.ascontiguousarray(params)
        assert params.dtype == render_params_dtype
        self._check(self.lib.nb_render_begin(self.handle, _ptr(params)))

    def render_subframe(self, first, count, stride=1, offset=0):
        self._check(self.lib.nb_render_subframe(self.handle, first, count, stride, offset))

    def adaptive_block_passes(self, n_blocks):
        out = np.zeros(n_blocks, "<u4")
        self._check(self.lib.nb_adaptive_block_passes(self.handle, _ptr(out), n_blocks))
        return out

    def render_adaptive_round(self, block_size=4, initial_uniform=2, round_cap=0):
        """One round of the adaptive sampler's feedback loop on the device (nb_render_adaptive_round):
        dict(samples, block_passes, passes, blocks)."""
        info = (C.c_uint64 * 3)()
        self._check(self.lib.nb_render_adaptive_round(self.handle, block_size, initial_uniform, round_cap, info))
        return {"samples": int(info[0]), "block_passes": int(info[1]), "passes": int(info[2] & 0xffffffff), "blocks": int(info[2] >> 32)}

    def sync(self):
        self._check(self.lib.nb_sync(self.handle))

    def film_clear(self):
        self._check(self.lib.nb_film_clear(self.handle))

    def film_shape(self):
        r, c, b = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._check(self.lib.nb_film_shape(self.handle, C.byref(r), C.byref(c), C.byref(b)))
        return r.value, c.value, b.value

    def film_download(self, normalize=True, out=None, layer=0):
        """layer 0: the composite; 1: the normal layer (needs render params with layers=1)."""
        rows, cols, b = self.film_shape()
        shape = (rows - 2 * b, cols - 2 * b, 3) if normalize else (rows, cols, 4)
        if out is None:
            out = np.empty(shape, dtype=np.float32)
        self._check(self.lib.nb_film_download_layer(self.handle, int(layer), _ptr(out), int(normalize)))
        return out

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_uint64()
        self._check(self.lib.nb_film_device_ptr(self.handle, C.byref(p), C.byref(n)))
        return p.value, n.value

    def counters(self):
        c = Counters()
        self._check(self.lib.nb_get_counters(self.handle, C.byref(c)))
        return {k: getattr(c, k) for k, _ in Counters._fields_}

    def set_option(self, name, value):
        self._check(self.lib.nb_set_option(self.handle, name.encode(), int(value)))


def make_render_params(scene, integrator=None, seed=0, pool_paths=0, max_depth=0, layers=0):
    """nb_render_params from a scene dict's "render" section (filter table, integrator) + overrides."""
    p = np.zeros(1, dtype=render_params_dtype)
    if "render" in scene:
        p[0] = scene["render"][0]
    if integrator is not None:
        p["integrator"] = INTEGRATORS[integrator] if isinstance(integrator, str) else int(integrator)
    elif "render" not in scene:
        p["integrator"] = INTEGRATORS["path_mis"]
    p["seed"] = seed
    p["pool_paths"] = pool_paths
    p["max_depth"] = max_depth
    p["layers"] = layers
    return p


def sample_slice(first, count, stride, offset):
    """(first_sample, n_samples) of one rank's slice — nb_sample_slice, host only."""
    a, b = C.c_uint32(), C.c_uint32()
    rc = load_library().nb_sample_slice(first, count, stride, offset, C.byref(a), C.byref(b))
    if rc != NB_OK:
        raise NbError(rc, "bad sample slice %d/%d" % (offset, stride))
    return a.value, b.value


# ---- multi-GPU: the library's own NCCL film reduce (include/norib200.h) ---------------------------
def comm_unique_id():
    """nb_comm_unique_id: the 128 bytes rank 0 hands to the other ranks (over the launcher's channel)."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    rc = load_library().nb_comm_unique_id(buf)
    if rc != NB_OK:
        raise NbError(rc, load_library().nb_last_error(None).decode())
    return buf.raw


def create_multi(devices):
    """nb_create_multi: one Context per device in THIS process, joined by one NCCL clique; [0] is the reduce root."""
    lib = load_library()
    devs = (C.c_int * len(devices))(*devices)
    out = (C.c_void_p * len(devices))()
    rc = lib.nb_create_multi(devs, len(devices), out)
    if rc != NB_OK:
        raise NbError(rc, lib.nb_last_error(None).decode())
    return [Context(d, _handle=h) for d, h in zip(devices, out)]


def comm_init_torch(ctx):
    """Join the library's communicator in a torchrun job: rank 0 makes the id, torch.distributed carries it."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    box = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ctx.comm_init(world, rank, box[0])


# ---- the same sum through torch.distributed (gloo in the CPU tests of the slice arithmetic) -------
def film_tensor(ctx, device_index):
    """torch tensor aliasing the context's film accumulators (no copy)."""
    import torch
    ptr, n = ctx.film_device_ptr()

    class _Alias:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_Alias(), device="cuda:%d" % device_index)


def reduce_film(film, dst=0):
    """The path's only collective: sum of the per-rank accumulators (sum w*rgb, sum w) on rank
    dst. Linear film (block.cpp:120-122) => the N-rank image is the 1-rank image of all samples."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
        if film.is_cuda:
            # the nb_* entries run on the library's own non-blocking stream, which never orders against torch's:
            # the host must see the collective finished before it touches the film through the library again
            import torch
            torch.cuda.current_stream(film.device).synchronize()
    return film


# ---- host library (C++ mirror of the Nori scene/plugin surface) ------------------------------------
_host = None


def load_host_library():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise FileNotFoundError("%s is not built: run `python optix-renderer_b200/build.py`" % HOST_LIB_PATH)
        load_library()                      # libnorib200_host.so links against libnorib200.so
        _host = C.CDLL(HOST_LIB_PATH)
        _host.nbh_last_error.restype = C.c_char_p
        _host.nbh_flatten_xml.argtypes = [C.c_char_p, C.c_char_p]
        _host.nbh_flatten_xml_opts.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        _host.nbh_parse_xml.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        _host.nbh_vdb_list.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        _host.nbh_vdb_read.restype = C.c_longlong
        _host.nbh_vdb_read.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        _host.nbh_vdb_to_nbv.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
        _host.nbh_to_nvdb.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        _host.nbh_nvdb_write_leaves.argtypes = [C.c_char_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        _host.nbh_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        _host.nbh_student_t_two_sided_p.restype = C.c_double; _host.nbh_student_t_two_sided_p.argtypes = [C.c_double, C.c_double]
        _host.nbh_chi2_upper_p.restype = C.c_double; _host.nbh_chi2_upper_p.argtypes = [C.c_double, C.c_double]
        _host.nbh_inflate.restype = C.c_longlong
        _host.nbh_inflate.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p, C.c_uint64]
    return _host


def vdb_list(path):
    """[(grid name, tree type)] of an OpenVDB file (host/vdb.cpp)."""
    h = load_host_library()
    buf = C.create_string_buffer(1 << 16)
    if h.nbh_vdb_list(os.fsencode(path), buf, len(buf)) != 0:
        raise NbError(-1, h.nbh_last_error().decode())
    return [tuple(l.split(":")) for l in buf.value.decode().splitlines()]


def vdb_read(path, grid="density"):
    """One float grid of an OpenVDB file as a dense array over its active box (host/vdb.cpp):
    dict(values (nx,ny,nz) f32, bbox_min, bbox_max, index_to_world (3,4), background, min, max, active_voxels)."""
    h = load_host_library()
    bbox = np.zeros(6, "<i4"); xf = np.zeros(12, "<f8"); st = np.zeros(3, "<f4"); act = C.c_uint64()
    n = h.nbh_vdb_read(os.fsencode(path), grid.encode(), _ptr(bbox), _ptr(xf), _ptr(st), C.byref(act), None, 0)
    if n < 0:
        raise NbError(-1, h.nbh_last_error().decode())
    v = np.zeros(n, "<f4")
    h.nbh_vdb_read(os.fsencode(path), grid.encode(), _ptr(bbox), _ptr(xf), _ptr(st), C.byref(act), _ptr(v), n)
    dims = bbox[3:] - bbox[:3] + 1
    return {"values": v.reshape(dims) if n else v, "bbox_min": bbox[:3].copy(), "bbox_max": bbox[3:].copy(), "index_to_world": xf.reshape(3, 4),
            "background": float(st[0]), "min": float(st[1]), "max": float(st[2]), "active_voxels": act.value}


def to_nvdb(src_path, nvdb_path, grid="density", sparse=False):
    """.vdb / .nbv / .nvdb -> the NanoVDB cache file the reference would write next to a .vdb (host/nvdb.cpp).
    sparse=True: only the voxels that differ from the background are active and only the 8^3 leaves that hold some are stored;
    sparse="file": the source file's own leaves and active masks, as openToNanoVDB keeps them (NvdbVolume.vdb.cpp:118-119)."""
    h = load_host_library()
    if h.nbh_to_nvdb(os.fsencode(src_path), grid.encode(), os.fsencode(nvdb_path), 2 if sparse == "file" else int(bool(sparse))) != 0:
        raise NbError(-1, h.nbh_last_error().decode())


def nvdb_write_leaves(nvdb_path, origins, values, index_to_world, grid="density", background=0.0):
    """A sparse float grid from 8^3 blocks — origins (n, 3) int32 multiples of 8, values (n, 8, 8, 8) f32, x slowest — as a
    .nvdb file (host/nvdb.cpp): voxels that differ from the background are active."""
    h = load_host_library()
    o = np.ascontiguousarray(origins, "<i4").reshape(-1, 3); v = np.ascontiguousarray(values, "<f4").reshape(len(o), 512)
    M = np.ascontiguousarray(index_to_world, "<f8").reshape(12)
    if h.nbh_nvdb_write_leaves(os.fsencode(nvdb_path), grid.encode(), len(o), _ptr(o), _ptr(v), _ptr(M), float(background)) != 0:
        raise NbError(-1, h.nbh_last_error().decode())


def load_xml(xml_path, nbs_path=None, host_bvh=True):
    """Scene XML -> flattened scene dict, through the C++ host loader (parser, OBJ loader, plugin
    flatten hooks, SAH BVH build). Raises NbError(NB_E_UNSUPPORTED) for plugins without a kernel.
    host_bvh=False skips the host SAH build (bvh_nodes / bvh_indices come back empty): load with
    Context.load(s, device_bvh="sah")."""
    import tempfile
    h = load_host_library()
    tmp = nbs_path or os.path.join(tempfile.mkdtemp(prefix="nbs_"), "scene.nbs")
    rc = h.nbh_flatten_xml_opts(os.fsencode(xml_path), os.fsencode(tmp), 1 if host_bvh else 0)
    if rc != 0:
        raise NbError(rc, h.nbh_last_error().decode())
    s = read_nbs(tmp)
    if nbs_path is None:
        os.remove(tmp)
    return s
