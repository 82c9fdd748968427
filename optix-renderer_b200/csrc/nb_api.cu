/* nb_api.cu — the C ABI of include/norib200.h: context, scene upload, BVH import, parity
 * trace entry. Rendering entries live in nb_render.cu. No CPU fallback anywhere: without a CUDA
 * device nb_create fails with NB_E_NO_DEVICE. */
#include "nb_internal.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>

static thread_local std::string g_create_error;

int nb_fail(nb_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->error = buf; else g_create_error = buf;
    return code;
}

template <typename T>
static int nb_upload(nb_ctx *ctx, std::vector<void *> &owner, const T *host, size_t count, const T **dev) {
    *dev = nullptr;
    if (count == 0 || host == nullptr) return NB_OK;
    void *p = nullptr;
    NB_CUDA(ctx, cudaMalloc(&p, count * sizeof(T)));
    owner.push_back(p);
    NB_CUDA(ctx, cudaMemcpyAsync(p, host, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *dev = (const T *) p;
    return NB_OK;
}

static void nb_free_list(std::vector<void *> &v) {
    for (void *p : v) cudaFree(p);
    v.clear();
}

extern "C" {

int nb_abi_version(void) { return NB_ABI_VERSION; }

int nb_abi_sizeof(int which) {
    switch (which) {
        case 0: return (int) sizeof(nb_shape);
        case 1: return (int) sizeof(nb_bsdf);
        case 2: return (int) sizeof(nb_emitter);
        case 3: return (int) sizeof(nb_camera);
        case 4: return (int) sizeof(nb_bvh_node);
        case 5: return (int) sizeof(nb_scene_desc);
        case 6: return (int) sizeof(nb_ray);
        case 7: return (int) sizeof(nb_hit);
        case 8: return (int) sizeof(nb_render_params);
        case 9: return (int) sizeof(nb_counters);
        case 10: return (int) sizeof(nb_medium);
        case 11: return (int) sizeof(nb_texture);
        default: return -1;
    }
}

const char *nb_last_error(const nb_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int nb_create(int device, nb_ctx **out) {
    if (!out) return nb_fail(nullptr, NB_E_INVALID, "nb_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return nb_fail(nullptr, NB_E_NO_DEVICE, "nb_create: no CUDA device (%s); this backend has no CPU fallback",
                       e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0 || device >= n) return nb_fail(nullptr, NB_E_INVALID, "nb_create: device %d out of range [0,%d)", device, n);
    nb_ctx *ctx = new nb_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete ctx;
        return nb_fail(nullptr, NB_E_CUDA, "nb_create: %s", cudaGetErrorString(e));
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = ctx;
    return NB_OK;
}

int nb_destroy(nb_ctx *ctx) {
    if (!ctx) return NB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    nb_comm_free(ctx);
    for (int i = 0; i < 2; ++i) if (ctx->timer[i]) cudaEventDestroy(ctx->timer[i]);
    nb_render_free(ctx);
    nb_free_list(ctx->scene_allocs);
    nb_free_list(ctx->bvh_allocs);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return NB_OK;
}

int nb_timer_mark(nb_ctx *ctx, int which) {
    if (!ctx || which < 0 || which > 1) return nb_fail(ctx, NB_E_INVALID, "nb_timer_mark: which must be 0 or 1");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->timer[which]) NB_CUDA(ctx, cudaEventCreate(&ctx->timer[which]));
    NB_CUDA(ctx, cudaEventRecord(ctx->timer[which], ctx->stream));
    return NB_OK;
}

int nb_timer_elapsed(nb_ctx *ctx, float *ms) {
    if (!ctx || !ms) return nb_fail(ctx, NB_E_INVALID, "nb_timer_elapsed: NULL argument");
    if (!ctx->timer[0] || !ctx->timer[1]) return nb_fail(ctx, NB_E_STATE, "nb_timer_elapsed: mark 0 and 1 first");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    NB_CUDA(ctx, cudaEventSynchronize(ctx->timer[1]));
    NB_CUDA(ctx, cudaEventElapsedTime(ms, ctx->timer[0], ctx->timer[1]));
    return NB_OK;
}

int nb_upload_scene(nb_ctx *ctx, const nb_scene_desc *s) {
    if (!ctx || !s) return nb_fail(ctx, NB_E_INVALID, "nb_upload_scene: NULL argument");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (s->n_shapes == 0 || !s->shapes) return nb_fail(ctx, NB_E_INVALID, "nb_upload_scene: scene has no shapes");
    /* validate before touching the device */
    uint64_t prims = 0;
    std::vector<int32_t> medium_emitter;
    bool ext_plugins = false;      /* disney / spot / directional: rendered by the EXT kernel variants */
    for (uint32_t i = 0; i < s->n_shapes; ++i) {
        const nb_shape &sh = s->shapes[i];
        if (sh.type == NB_SHAPE_MESH) {
            if ((uint64_t) sh.vertex_offset + sh.vertex_count > s->n_vertices || (uint64_t) sh.face_offset + sh.prim_count > s->n_faces)
                return nb_fail(ctx, NB_E_INVALID, "shape %u: vertex/face range outside the arrays", i);
            for (uint32_t f = 0; f < sh.prim_count * 3u; ++f)
                if (s->faces[(size_t) sh.face_offset * 3 + f] >= sh.vertex_count)
                    return nb_fail(ctx, NB_E_INVALID, "shape %u: face index out of range", i);
            if ((sh.flags & NB_SHAPE_HAS_NORMALS) && !s->normals) return nb_fail(ctx, NB_E_INVALID, "shape %u: normals flagged but absent", i);
            if ((sh.flags & NB_SHAPE_HAS_UVS) && !s->uvs) return nb_fail(ctx, NB_E_INVALID, "shape %u: uvs flagged but absent", i);
            if (sh.emitter >= 0 && (uint64_t) sh.cdf_offset + sh.prim_count + 1 > s->n_area_cdf)
                return nb_fail(ctx, NB_E_INVALID, "shape %u: area CDF range outside the array", i);
        } else if (sh.type == NB_SHAPE_SPHERE) {
            if (sh.prim_count != 1) return nb_fail(ctx, NB_E_INVALID, "shape %u: a sphere has exactly one primitive", i);
        } else return nb_fail(ctx, NB_E_UNSUPPORTED, "shape %u: unknown shape tag %u", i, sh.type);
        if (sh.bsdf >= (int32_t) s->n_bsdfs || sh.emitter >= (int32_t) s->n_emitters || sh.medium >= (int32_t) s->n_media)
            return nb_fail(ctx, NB_E_INVALID, "shape %u: bsdf/emitter/medium index out of range", i);
        prims += sh.prim_count;
    }
    if (prims >= NB_SLOT_LAST) return nb_fail(ctx, NB_E_UNSUPPORTED, "too many primitives (%llu)", (unsigned long long) prims);
    for (uint32_t i = 0; i < s->n_bsdfs; ++i) {
        const uint32_t tag = s->bsdfs[i].type & 0xffu, tex = s->bsdfs[i].type >> 8;
        if (tag > NB_BSDF_DISNEY) return nb_fail(ctx, NB_E_UNSUPPORTED, "bsdf %u: tag %u has no kernel in this backend", i, tag);
        if (tex && tag != NB_BSDF_DIFFUSE && tag != NB_BSDF_DISNEY)
            return nb_fail(ctx, NB_E_UNSUPPORTED, "bsdf %u: only the diffuse and disney BSDFs take an albedo texture", i);
        if (tex > s->n_textures) return nb_fail(ctx, NB_E_INVALID, "bsdf %u: texture index out of range", i);
        if (tex && s->textures && (s->textures[tex - 1].flags & NB_TEXTURE_NORMAL_MAP)) return nb_fail(ctx, NB_E_INVALID, "bsdf %u: albedo texture is a normal map", i);
        if (tag == NB_BSDF_DISNEY) {
            uint32_t off; memcpy(&off, &s->bsdfs[i].alpha, 4);
            if (!s->params || (uint64_t) off + 10 > s->n_params) return nb_fail(ctx, NB_E_INVALID, "bsdf %u: disney parameters outside params", i);
            ext_plugins = true;
        }
    }
    for (uint32_t i = 0; i < s->n_shapes; ++i) {
        const uint32_t nm = s->shapes[i].normal_map;
        if (!nm) continue;
        if (nm > s->n_textures || !s->textures || s->textures[nm - 1].type != NB_TEXTURE_IMAGE || !(s->textures[nm - 1].flags & NB_TEXTURE_NORMAL_MAP))
            return nb_fail(ctx, NB_E_INVALID, "shape %u: normal_map does not name a normal-map texture", i);
    }
    if (s->n_textures && !s->textures) return nb_fail(ctx, NB_E_INVALID, "textures missing");
    for (uint32_t i = 0; i < s->n_textures; ++i) {
        const nb_texture &t = s->textures[i];
        if (t.type > NB_TEXTURE_CHECKERBOARD) return nb_fail(ctx, NB_E_UNSUPPORTED, "texture %u: tag %u has no kernel in this backend", i, t.type);
        if (t.width == 0 || t.height == 0) return nb_fail(ctx, NB_E_INVALID, "texture %u: empty", i);
        if (t.type == NB_TEXTURE_IMAGE && (!s->texels || (uint64_t) t.first_texel + (uint64_t) t.width * t.height > s->n_texels))
            return nb_fail(ctx, NB_E_INVALID, "texture %u: texels outside the texel array", i);
        if (t.type == NB_TEXTURE_CHECKERBOARD && (t.scale[0] == 0.f || t.scale[1] == 0.f))
            return nb_fail(ctx, NB_E_INVALID, "texture %u: checkerboard scale is zero", i);
    }
    int32_t envmap = -1;
    for (uint32_t i = 0; i < s->n_emitters; ++i) {
        const nb_emitter &e = s->emitters[i];
        if (e.type == NB_EMITTER_AREA) {
            if (e.shape < 0 || e.shape >= (int32_t) s->n_shapes) return nb_fail(ctx, NB_E_INVALID, "emitter %u: area light without a shape", i);
        } else if (e.type == NB_EMITTER_POINT) {
        } else if (e.type == NB_EMITTER_VOLUME) {
            uint32_t off; memcpy(&off, &e.position[1], 4);
            if (e.shape < 0 || e.shape >= (int32_t) s->n_shapes) return nb_fail(ctx, NB_E_INVALID, "emitter %u: volume light without a shape", i);
            if (s->shapes[e.shape].type == NB_SHAPE_MESH && (!s->params || (uint64_t) off + 6 > s->n_params))
                return nb_fail(ctx, NB_E_INVALID, "emitter %u: volume light box outside params", i);
            if (!(e.position[0] > 0.f)) return nb_fail(ctx, NB_E_INVALID, "emitter %u: volume light needs 1 / volume > 0", i);
            if (s->shapes[e.shape].emitter != (int32_t) i) {          /* the light of the shape's medium */
                const int32_t m = s->shapes[e.shape].medium;
                if (m < 0) return nb_fail(ctx, NB_E_INVALID, "emitter %u: volume light on a shape without a medium", i);
                medium_emitter.resize(s->n_media, -1);
                medium_emitter[m] = (int32_t) i;
            }
            ext_plugins = true;
        } else if (e.type == NB_EMITTER_SPOT || e.type == NB_EMITTER_DIRECTIONAL) {
            const uint64_t need = e.type == NB_EMITTER_SPOT ? 5 : 4;
            if (e.shape < 0 || !s->params || (uint64_t) e.shape + need > s->n_params)
                return nb_fail(ctx, NB_E_INVALID, "emitter %u: parameters outside params", i);
            ext_plugins = true;
        } else if (e.type == NB_EMITTER_ENVMAP) {                  /* Scene::addChild keeps the last one as m_envmap (scene.cpp:218-222) */
            if (e.shape < 0 || e.shape >= (int32_t) s->n_textures) return nb_fail(ctx, NB_E_INVALID, "emitter %u: environment map without a texture", i);
            const nb_texture &t = s->textures[e.shape];
            uint32_t off; memcpy(&off, &e.position[0], 4);
            if (!(t.width == 1 && t.height == 1) && (uint64_t) off + (uint64_t) t.width * t.height + 1 > s->n_area_cdf)
                return nb_fail(ctx, NB_E_INVALID, "emitter %u: environment map CDF outside area_cdf", i);
            envmap = (int32_t) i;
        } else return nb_fail(ctx, NB_E_UNSUPPORTED, "emitter %u: tag %u has no kernel in this backend", i, e.type);
    }
    if (s->n_emitters && !s->emitter_cdf) return nb_fail(ctx, NB_E_INVALID, "emitter_cdf missing");
    if (s->n_media && !s->media) return nb_fail(ctx, NB_E_INVALID, "media missing");
    if (s->ambient_medium >= (int32_t) s->n_media) return nb_fail(ctx, NB_E_INVALID, "ambient medium index out of range");
    for (uint32_t i = 0; i < s->n_media; ++i) {
        const nb_medium &m = s->media[i];
        if (m.type > NB_MEDIUM_HETEROGENEOUS_BRICKS || m.phase > NB_PHASE_SCHLICK)
            return nb_fail(ctx, NB_E_UNSUPPORTED, "medium %u: medium tag %u / phase tag %u has no kernel in this backend", i, m.type, m.phase);
        if (m.type >= NB_MEDIUM_HETEROGENEOUS) {
            const uint64_t vox = (uint64_t) m.grid_dims[0] * m.grid_dims[1] * m.grid_dims[2];
            if (m.type == NB_MEDIUM_HETEROGENEOUS_BRICKS) {
                /* sparse layout: the table, then every brick a table entry names, inside the density array */
                const uint64_t entries = (uint64_t) ((m.grid_dims[0] + 7) / 8) * ((m.grid_dims[1] + 7) / 8) * ((m.grid_dims[2] + 7) / 8);
                if (vox == 0 || !s->density || (uint64_t) m.first_voxel + entries + NB_BRICK_VOXELS > s->n_density)
                    return nb_fail(ctx, NB_E_INVALID, "medium %u: brick table outside the density array", i);
                const uint64_t room = (s->n_density - m.first_voxel - entries) / NB_BRICK_VOXELS;
                for (uint64_t e = 0; e < entries; ++e) {
                    uint32_t brick; memcpy(&brick, s->density + m.first_voxel + e, 4);
                    if (brick >= room) return nb_fail(ctx, NB_E_INVALID, "medium %u: brick table entry %llu names brick %u of %llu", i, (unsigned long long) e, brick, (unsigned long long) room);
                }
            } else if (vox == 0 || !s->density || (uint64_t) m.first_voxel + vox > s->n_density)
                return nb_fail(ctx, NB_E_INVALID, "medium %u: density grid outside the density array", i);
            for (int a = 0; a < 3; ++a)
                if (m.grid_dims[a] < 2 * NB_GRID_BORDER + 1)
                    return nb_fail(ctx, NB_E_INVALID, "medium %u: grid needs a border of %d voxels and at least one voxel inside", i, NB_GRID_BORDER);
            if (!(m.max_density > 0.f) || !(fmaxf(m.mu_t[0], fmaxf(m.mu_t[1], m.mu_t[2])) > 0.f))
                return nb_fail(ctx, NB_E_INVALID, "medium %u: heterogeneous medium needs max_density > 0 and sigma_t > 0", i);
        }
    }

    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }     /* queued subframes still read the old scene */
    nb_free_list(ctx->scene_allocs);
    nb_free_list(ctx->bvh_allocs);
    ctx->scene_ready = ctx->bvh_ready = ctx->render_begun = false;
    ctx->dev = NbDeviceScene();
    ctx->dev.ordered = ctx->opt_ordered ? 1u : 0u;

    ctx->h_shapes.assign(s->shapes, s->shapes + s->n_shapes);
    ctx->h_positions.assign(s->positions, s->positions + s->n_vertices * 3);
    ctx->h_faces.assign(s->faces, s->faces + s->n_faces * 3);
    ctx->h_shape_prim_offset.assign(s->n_shapes + 1, 0);
    for (uint32_t i = 0; i < s->n_shapes; ++i) ctx->h_shape_prim_offset[i + 1] = ctx->h_shape_prim_offset[i] + s->shapes[i].prim_count;

    NbDeviceScene &d = ctx->dev;
    int rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->shapes, s->n_shapes, &d.shapes))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->positions, s->n_vertices * 3, &d.positions))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->normals, s->n_vertices * 3, &d.normals))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->uvs, s->n_vertices * 2, &d.uvs))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->tangents, s->n_vertices * 3, &d.tangents))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->bitangents, s->n_vertices * 3, &d.bitangents))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->faces, s->n_faces * 3, &d.faces))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->bsdfs, s->n_bsdfs, &d.bsdfs))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->emitters, s->n_emitters, &d.emitters))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->emitter_cdf, s->n_emitters ? s->n_emitters + 1 : 0, &d.emitter_cdf))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->area_cdf, s->n_area_cdf, &d.area_cdf))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->media, s->n_media, &d.media))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->density, s->n_density, &d.density))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->textures, s->n_textures, &d.textures))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->texels, s->n_texels * 4, &d.texels))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, s->params, s->n_params, &d.params))) return rc;
    if ((rc = nb_upload(ctx, ctx->scene_allocs, medium_emitter.data(), medium_emitter.size(), &d.medium_emitter))) return rc;
    d.n_media = s->n_media; d.ambient_medium = s->ambient_medium;
    d.n_shapes = s->n_shapes; d.n_bsdfs = s->n_bsdfs; d.n_emitters = s->n_emitters;
    d.envmap_emitter = envmap;
    d.ext = (envmap >= 0 || s->n_textures > 0 || ext_plugins) ? 1u : 0u;
    for (uint32_t i = 0; i < s->n_media; ++i) if (s->media[i].type >= NB_MEDIUM_HETEROGENEOUS) d.ext = 1u;
    d.camera = s->camera;
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->scene_ready = true;
    return NB_OK;
}

} /* extern "C" */
