/* norib200.h — C ABI of the B200-native path-tracing backend (libnorib200.so).
 *
 * This is the drop-in boundary for the reference's GPU side-car. It replaces the call shape of
 * `OptixState` (reference include/nori/optix/OptixState.h:88-109: preRender / renderSubframe /
 * clear) and of the flatten hooks that feed it (getOptixMaterialData bsdf.h:122,
 * getOptixEmitterData emitter.h:159, getOptixData camera.h:86, getOptixBuildInput shape.h:194-201,
 * gathered in OptixState::preRender, OptixState.render.cpp:19-85). The caller is
 * RenderThread::renderThreadOptix (src/utils/render.cpp:462-574).
 *
 * Conventions: plain C, plain pointers and sizes, no C++/torch types. Every entry returns an
 * int (NB_OK == 0, negative NB_E_*), never throws or aborts across the boundary;
 * nb_last_error() returns the text of the last failure on that context (or the global one for
 * nb_create). All `const T*` inputs are HOST pointers copied during the call (the caller keeps
 * ownership) unless the name says `_device`.
 */
#ifndef NORIB200_H
#define NORIB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_ABI_VERSION 7

enum {
    NB_OK = 0,
    NB_E_INVALID = -1,      /* bad argument / inconsistent scene description */
    NB_E_CUDA = -2,         /* CUDA runtime error (text in nb_last_error) */
    NB_E_NOMEM = -3,
    NB_E_STATE = -4,        /* call out of order (e.g. render before upload) */
    NB_E_UNSUPPORTED = -5,  /* a plugin tag this backend has no kernel for: caller renders on CPU
                               (mirrors preRender() == false, render.cpp:488-489) */
    NB_E_NO_DEVICE = -6     /* no CUDA device: there is NO CPU fallback in this library */
};

/* ---- plugin tags (the flatten hooks emit these) -------------------------------------------- */
enum { NB_SHAPE_MESH = 0, NB_SHAPE_SPHERE = 1 };
enum { NB_BSDF_DIFFUSE = 0, NB_BSDF_MIRROR = 1, NB_BSDF_DIELECTRIC = 2, NB_BSDF_MICROFACET = 3,
       NB_BSDF_DISNEY = 4, NB_BSDF_NONE = 0xff };
enum { NB_EMITTER_AREA = 0, NB_EMITTER_POINT = 1, NB_EMITTER_ENVMAP = 2, NB_EMITTER_SPOT = 3,
       NB_EMITTER_DIRECTIONAL = 4, NB_EMITTER_VOLUME = 5 };
enum { NB_TEXTURE_CONSTANT = 0, NB_TEXTURE_IMAGE = 1, NB_TEXTURE_CHECKERBOARD = 2 };
enum { NB_TEXTURE_SPHERICAL = 1u,   /* PNGTexture::sphericalTexture */
       NB_TEXTURE_NORMAL_MAP = 2u }; /* PNGTexture with sRGB = false: a Texture<Normal3f> (PNGTexture.cpp:88-99, 155-162) */
enum { NB_INTEGRATOR_PATH_MIS = 0, NB_INTEGRATOR_PATH_MATS = 1, NB_INTEGRATOR_DIRECT_EMS = 2,
       NB_INTEGRATOR_DIRECT_MATS = 3, NB_INTEGRATOR_DIRECT_MIS = 4, NB_INTEGRATOR_NORMALS = 5,
       NB_INTEGRATOR_DIRECT = 6, NB_INTEGRATOR_AV = 7, NB_INTEGRATOR_PATH_VOL_MATS = 8,
       NB_INTEGRATOR_PATH_VOL_MIS = 9, NB_INTEGRATOR_PREVIEW = 10 };
enum { NB_MEDIUM_VACUUM = 0, NB_MEDIUM_HOMOGENEOUS = 1, NB_MEDIUM_HETEROGENEOUS = 2, NB_MEDIUM_HETEROGENEOUS_BRICKS = 3 };
enum { NB_PHASE_ISOTROPIC = 0, NB_PHASE_HENYEY_GREENSTEIN = 1, NB_PHASE_SCHLICK = 2 };
enum { NB_SHAPE_HAS_NORMALS = 1u, NB_SHAPE_HAS_UVS = 2u };

/* ---- scene records ------------------------------------------------------------------------- */

/* One Shape (src/shapes/mesh.cpp, sphere.cpp; include/nori/shape.h:104-237). Primitive numbering
 * is the BVH's: shapes concatenated in scene order (BVH::addShape, bvh.cpp:308-317). */
typedef struct nb_shape {
    uint32_t type;           /* NB_SHAPE_* */
    uint32_t prim_count;     /* triangles, or 1 for a sphere */
    uint32_t vertex_offset;  /* first vertex of this mesh in positions/normals/uvs */
    uint32_t vertex_count;
    uint32_t face_offset;    /* first triangle of this mesh in faces (3 u32 each, mesh-local ids) */
    uint32_t flags;          /* NB_SHAPE_HAS_* */
    int32_t  bsdf;           /* index into bsdfs, -1 = none (medium interface) */
    int32_t  emitter;        /* index into emitters, -1 = not a light */
    int32_t  medium;         /* index into media (Shape::getMedium, shape.h:146), -1 = none */
    uint32_t cdf_offset;     /* first of prim_count+1 floats in area_cdf (DiscretePDF::m_cdf) */
    float    pdf_norm;       /* DiscretePDF::getNormalization() = 1/total area (mesh.cpp:70) */
    float    radius;         /* sphere */
    float    center[3];      /* sphere */
    uint32_t normal_map;     /* 1 + index of the NB_TEXTURE_NORMAL_MAP texture (Shape::m_normalMap, shape.h:226;
                                applied in mesh.cpp:168-186, sphere.cpp:115-123), 0 = none */
} nb_shape;                  /* 64 bytes */

/* BSDF::getOptixMaterialData replacement (bsdf/diffuse.cpp, mirror.cpp, dielectric.cpp,
 * microfacet.cpp). */
typedef struct nb_bsdf {
    uint32_t type;           /* bits 0..7: NB_BSDF_*; bits 8..31: 1 + index of the albedo texture
                                (Diffuse::m_albedo, diffuse.cpp:31-40,77-91; Disney::albedo, disney.cpp:22-29,
                                90-104), 0 = the constant below */
    float    albedo[3];      /* diffuse albedo / microfacet kd / disney base colour */
    float    alpha;          /* microfacet roughness. Disney (src/bsdf/disney.cpp:31-41): the bits of a u32 —
                                offset into nb_scene_desc.params of its 10 floats { metallic, subsurface,
                                specular, roughness, specularTint, anisotropic, sheen, sheenTint, clearcoat,
                                clearcoatGloss }, each already clamped to [0, 1] */
    float    int_ior, ext_ior;
    float    ks;             /* microfacet: 1 - max(kd) (microfacet.cpp:52) */
} nb_bsdf;                   /* 32 bytes */

/* Emitter::getOptixEmitterData replacement (emitters/arealight.cpp, pointlight.cpp,
 * environmentmap.cpp). */
typedef struct nb_emitter {
    uint32_t type;           /* NB_EMITTER_* */
    int32_t  shape;          /* area light: owning shape; environment map: index of its texture (EnvMap::m_map);
                                spot / directional: offset into nb_scene_desc.params —
                                spot (src/emitters/spotlight.cpp:16-24, 63-70, 167-187): { direction.xyz (normalized),
                                  cos(falloffStart deg), cos(totalWidth / 2 deg) }, position below, radiance = power / (2 pi);
                                directional (src/emitters/directionalLight.cpp:31-38, 90-133): { direction.xyz
                                  (normalized), cos(angle deg) }, radiance below;
                                volume light (src/emitters/volumelight.cpp:47-77): the shape whose volume emits: the
                                  shape's own emitter when nb_shape.emitter names this record, else the emitter of
                                  its medium (Medium::getEmitter, medium.cpp:59-64); the three position words are
                                  { f32 1 / volume (Shape::pdfVolume, shape.cpp:103-111; sphere.h:45), u32 offset into
                                  params of { bbox min.xyz, bbox extents.xyz } (Shape::sampleVolume, shape.cpp:97-101), 0 } */
    float    radiance[3];    /* area: radiance; point: power/(4 pi) (pointlight.cpp:44); envmap: m_radiance */
    float    position[3];    /* point light. Environment map: the three words are
                                { u32 cdf_offset, f32 cdf_norm, 0 } — EnvMap::dpdf (environmentmap.cpp:146-161):
                                width*height+1 floats at cdf_offset of area_cdf, and its getNormalization() */
} nb_emitter;                /* 32 bytes */

/* Texture<Color3f> (src/textures/consttexture.cpp:37, checkerboard.cpp:34-52, PNGTexture.cpp:125-165),
 * replaces Texture::getOptixTexture (texture.h:75). An image's texels are RGBA floats exactly as
 * PNGTexture::data holds them after loadFromFile (inverse gamma applied), row 0 first. */
typedef struct nb_texture {
    uint32_t type;           /* NB_TEXTURE_* */
    uint32_t flags;          /* NB_TEXTURE_SPHERICAL: PNGTexture::sphericalTexture */
    uint32_t width, height;  /* Texture::getWidth/getHeight (1 x 1 for constant and checkerboard) */
    uint32_t first_texel;    /* image: offset into nb_scene_desc.texels, in texels (4 floats each) */
    float    value[3];       /* constant: the colour; checkerboard: value1 */
    float    value2[3];      /* checkerboard: value2 */
    float    scale[2];       /* image: scaleU, scaleV; checkerboard: m_scale */
    float    offset[2];      /* image: offsetU, offsetV; checkerboard: m_delta */
    float    rotation[9];    /* image, spherical: the row-major rotation PNGTexture::eval builds from eulerAngles */
    float    intensity;      /* normal map: PNGTexture::intensity (PNGTexture.cpp:33, 155-162) */
    uint32_t pad;
} nb_texture;                /* 96 bytes */

/* Camera::getOptixData replacement (cameras/perspective.cpp:48-141). Matrices are row-major
 * 4x4 as Eigen prints them (element [r*4+c]). */
typedef struct nb_camera {
    float    sample_to_camera[16];
    float    camera_to_world[16];
    float    inv_output_size[2];
    float    near_clip, far_clip;
    float    lens_radius, focal_distance;
    uint32_t width, height;
} nb_camera;                 /* 160 bytes */

/* BVHNode of the reference (include/nori/bvh.h:127-164), bit for bit: word0 bit0 = leaf flag,
 * bits 1..31 = leaf size | split axis; word1 = leaf start | right child; then bbox min, max. */
typedef struct nb_bvh_node {
    uint32_t word0, word1;
    float    bmin[3], bmax[3];
} nb_bvh_node;               /* 32 bytes */

/* One Medium with its phase function (replaces Medium::getOptixMediumData, medium.h:68).
 * Homogeneous (src/media/homogmedium.cpp:48-73): mu_t = density * (sigma_a + sigma_s) per channel.
 * Heterogeneous (src/media/heterogmedium.cpp:67-174): the density grid the reference samples through
 * NanoVDB, as a dense array: grid_dims voxels (x slowest, z fastest) at first_voxel of
 * nb_scene_desc.density, holding the box of active voxels (NanoGrid::indexBBox) plus a border of
 * NB_GRID_BORDER background voxels on every side. world_to_grid maps a world point to array
 * coordinates (voxel (i,j,k) of the array has its value at (i,j,k), as NanoVDB's index space), so
 * the medium's bounding box is [NB_GRID_BORDER, grid_dims - NB_GRID_BORDER] per axis.
 * max_density = the unscaled grid maximum (tree().root().valueMax(), heterogmedium.cpp:79).
 * NB_MEDIUM_HETEROGENEOUS_BRICKS: the same medium over a SPARSE grid (what NanoVDB's tree is for; a 1024^3 box of smoke
 * would be 4 GB dense). grid_dims, world_to_grid and the border mean what they mean above — the array is virtual. At
 * first_voxel the density array holds a table of ceil(grid_dims / 8) entries per axis (x slowest), each the bit pattern of
 * a uint32 brick number, followed by the bricks: NB_BRICK_VOXELS floats each, the 8^3 voxels of the block plus one voxel
 * of its neighbours on every side (10^3, x slowest), so the 27-tap stencil of any point in a block reads ONE brick.
 * Brick 0 is all background; table entries of blocks that hold nothing else point to it. */
#define NB_GRID_BORDER 2
#define NB_BRICK_EDGE 10
#define NB_BRICK_VOXELS 1000
typedef struct nb_medium {
    uint32_t type;           /* NB_MEDIUM_* */
    float    mu_t[3];        /* homogeneous: extinction; heterogeneous: sigma_t = sigma_a + sigma_s */
    uint32_t phase;          /* NB_PHASE_* (src/bsdf/isophase.cpp, anisophase.cpp, schlickphase.cpp) */
    float    phase_param;    /* g (Henyey-Greenstein) or k = 1.55 g - 0.55 g^3 (Schlick) */
    float    density_scale;  /* heterogeneous */
    float    max_density;    /* heterogeneous: unscaled grid maximum */
    uint32_t grid_dims[3];   /* heterogeneous: voxels in x, y, z */
    uint32_t first_voxel;    /* heterogeneous: offset into nb_scene_desc.density */
    float    world_to_grid[12]; /* heterogeneous: row-major 3x4, world point -> voxel coordinates */
} nb_medium;                 /* 96 bytes */

typedef struct nb_scene_desc {
    uint32_t n_shapes;    const nb_shape   *shapes;
    uint64_t n_vertices;  const float      *positions;   /* 3 per vertex, world space */
                          const float      *normals;     /* 3 per vertex or NULL */
                          const float      *uvs;         /* 2 per vertex or NULL */
                          const float      *tangents;    /* 3 per vertex or NULL (Mesh::m_T)  */
                          const float      *bitangents;  /* 3 per vertex or NULL (Mesh::m_BT) */
    uint64_t n_faces;     const uint32_t   *faces;       /* 3 per triangle */
    uint32_t n_bsdfs;     const nb_bsdf    *bsdfs;
    uint32_t n_emitters;  const nb_emitter *emitters;
                          const float      *emitter_cdf; /* n_emitters+1 (scene.cpp:180-184) */
    uint64_t n_area_cdf;  const float      *area_cdf;
    nb_camera camera;
    uint32_t n_media;     const nb_medium  *media;       /* may be 0 / NULL */
    int32_t  ambient_medium;                             /* Scene::getAmbientMedium, -1 = vacuum */
    uint64_t n_density;   const float      *density;     /* voxel data of heterogeneous media or NULL */
    uint32_t n_textures;  const nb_texture *textures;    /* may be 0 / NULL */
    uint64_t n_texels;    const float      *texels;      /* 4 floats per texel, or NULL */
    uint64_t n_params;    const float      *params;      /* parameters of plugins whose record has no room for
                                                            them (disney, spot, directional), or NULL */
} nb_scene_desc;

/* nb_build_bvh(mode = NB_BVH_IMPORT): the caller's own arrays (BVH::m_nodes / m_indices after
 * Scene::update, scene.cpp:187-194). Same tree, same visiting order => bit-identical hits. */
enum { NB_BVH_IMPORT = 0,   /* the caller's tree (the reference's, or libnorib200_host's restatement of it) */
       NB_BVH_DEVICE = 1,   /* K_bvh_build: linear BVH built on the GPU from the uploaded scene; `imp` is ignored */
       NB_BVH_DEVICE_SAH = 2 }; /* the reference's SAH criterion (bvh.cpp:100-233) built top-down on the GPU, one tree level per pass:
                             * the quality of the host build in milliseconds instead of seconds; `imp` is ignored */
typedef struct nb_bvh_import {
    uint32_t n_nodes;    const nb_bvh_node *nodes;
    uint32_t n_indices;  const uint32_t    *indices;    /* global primitive ids, leaf order */
} nb_bvh_import;

/* ---- rays (parity entry) ------------------------------------------------------------------- */
typedef struct nb_ray { float o[3], d[3], mint, maxt; } nb_ray;      /* 32 bytes, Ray3f fields */
typedef struct nb_hit {
    float   t, u, v;     /* as Shape::rayIntersect returns them (before setHitInformation) */
    int32_t shape;       /* index into shapes, -1 = miss */
    int32_t prim;        /* primitive index local to the shape */
    uint32_t any_hit;    /* result of the shadow-ray query on the same ray (any_hit mode) */
} nb_hit;                /* 24 bytes */

/* ---- rendering ----------------------------------------------------------------------------- */
#define NB_FILTER_RESOLUTION 32   /* NORI_FILTER_RESOLUTION, include/nori/rfilter.h:25 */

/* Render layers of the reference (integrator.h:29-39). Composite is the film itself. Normal: the shading normal of
 * the first hit as PreviewIntegrator reports it (PreviewIntegrator.cpp:37), 0 from every other integrator and on a
 * miss; like every ImageBlock it drops samples with a negative component (block.cpp:94). Albedo: 0 in every
 * integrator of the reference, so it has no buffer here. */
enum { NB_LAYER_COMPOSITE = 0, NB_LAYER_NORMAL = 1 };

typedef struct nb_render_params {
    uint32_t integrator;                       /* NB_INTEGRATOR_* */
    float    filter_radius;                    /* ReconstructionFilter::getRadius() */
    float    filter_table[NB_FILTER_RESOLUTION + 1]; /* ImageBlock::m_filter (block.cpp:58-64) */
    uint64_t seed;                             /* mixed into every per-pixel pcg32 stream */
    uint32_t pool_paths;                       /* path slots in flight; 0 = library default */
    uint32_t max_depth;                        /* 0 = unbounded, as the reference */
    float    integrator_params[3];             /* normals: NormalIntegrator::direction (normals.cpp:12, default 0 0 1);
                                                  av: [0] = AverageVisibilityIntegrator::m_length (av.cpp:14) */
    uint32_t layers;                           /* NB_LAYER_* bits: extra render layers to accumulate beside the composite
                                                  (ERenderLayer, integrator.h:29-39; renderBlock puts them at render.cpp:455-457) */
} nb_render_params;

typedef struct nb_counters {
    uint64_t samples;        /* camera paths started (renderBlock iterations, render.cpp:436) */
    uint64_t rays_closest;   /* BVH::rayIntersect(shadowRay=false) queries, probe/main deduped */
    uint64_t rays_any;       /* BVH::rayIntersect(shadowRay=true) queries */
    uint64_t iterations;     /* wavefront iterations launched */
    uint64_t kernel_launches;
    uint64_t invalid_samples;/* dropped by the isValid() rule of ImageBlock::put (block.cpp:94) */
    double   device_ms;      /* CUDA-event time of the last nb_sync()ed render span */
    double   trace_ms, shade_ms, regen_ms; /* per-kernel sums when the "profile" option is on */
} nb_counters;

typedef struct nb_ctx nb_ctx;

int  nb_abi_version(void);
/* sizeof() of the records above as this build sees them; which = 0 shape, 1 bsdf, 2 emitter,
 * 3 camera, 4 bvh_node, 5 scene_desc, 6 ray, 7 hit, 8 render_params, 9 counters */
int  nb_abi_sizeof(int which);

int  nb_create(int device, nb_ctx **out);       /* fails with NB_E_NO_DEVICE when there is none */
/* One context per listed device, in ONE process, joined by one NCCL communicator clique (ncclCommInitAll): the shape
 * the reference's single render thread needs (src/utils/render.cpp:462-574 drives one OptixState). out gets n
 * contexts; out[0] is the root of nb_film_reduce. Upload the scene / build the BVH / nb_render_begin on each, give
 * context g the sample slice (stride n, offset g) in nb_render_subframe (asynchronous, so the loop over contexts keeps
 * all devices busy), then nb_film_reduce(out[0]) and nb_film_download(out[0]). Destroy each with nb_destroy. */
int  nb_create_multi(const int *devices, int n, nb_ctx **out);
int  nb_destroy(nb_ctx *ctx);
const char *nb_last_error(const nb_ctx *ctx);   /* ctx may be NULL: error of the last nb_create */

/* OptixState::preRender (OptixState.render.cpp:19-85): flatten + upload. */
int  nb_upload_scene(nb_ctx *ctx, const nb_scene_desc *scene);
int  nb_build_bvh(nb_ctx *ctx, int mode, const nb_bvh_import *import);
/* node / primitive-slot counts of the current tree and the device time of the last NB_BVH_DEVICE
 * build (0 for an imported tree); any pointer may be NULL */
int  nb_bvh_stats(nb_ctx *ctx, uint32_t *n_nodes, uint32_t *n_slots, double *build_ms);

/* Parity entry: Scene::rayIntersect (include/nori/scene.h:114-138) over n host rays.
 * any_hit = 0: closest hit per ray. any_hit = 1: also runs the shadow-ray query and fills
 * nb_hit.any_hit. */
int  nb_trace_rays(nb_ctx *ctx, const nb_ray *rays, uint64_t n, int any_hit, nb_hit *out);
/* Same on rays already resident on the device (timing the kernel alone); out_device gets
 * n float4 {t,u,v,bits(prim slot)}; returns the kernel's CUDA-event time. */
int  nb_trace_rays_device(nb_ctx *ctx, const void *rays_device, uint64_t n, int any_hit,
                          void *out_device, int repeats, float *ms_per_launch);
/* nb_trace_rays for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). It has no counterpart in the reference.
 * Same arguments, same results bit for bit as nb_trace_rays, but it calls no cudaMalloc / cudaFree / allocating
 * library routine and no host heap: the device staging is a fixed file-scope __device__ array inside the library
 * (3.5 MiB, held by every GPU that loads the library until exit), so n is limited to NB_TRACE_STATIC_MAX_RAYS —
 * a larger n fails with NB_E_INVALID and a message saying so. All calls on one GPU share that array (it keeps the
 * last call's contents; nothing in it needs clearing): calls that overlap on the same GPU race and must be
 * serialised by the caller. Synchronous: `out` is complete on return. */
#define NB_TRACE_STATIC_MAX_RAYS 65536u
int  nb_trace_rays_static(nb_ctx *ctx, const nb_ray *rays, uint64_t n, int any_hit, nb_hit *out);
/* Parity entry for the medium rows: Medium::sampleFreePath (src/media/homogmedium.cpp:60-66,
 * heterogmedium.cpp:105-130), Medium::getTransmittance (homogmedium.cpp:68-72, heterogmedium.cpp:132-174)
 * and NvdbVolume::getDensity (src/textures/NvdbVolume.cpp:39-48) of media[medium] on n host rays.
 * out gets 4 floats per ray: density at ray.o (0 for a medium without a grid); the fraction of
 * `samples` free paths from ray.o along ray.d shorter than ray.maxt; their mean length; the mean
 * transmittance (first channel) from ray.o to ray.o + ray.maxt * ray.d. */
int  nb_medium_probe(nb_ctx *ctx, int32_t medium, const nb_ray *rays, uint64_t n, uint32_t samples,
                     uint64_t seed, float *out);
/* Parity entry for the emitter rows: Emitter::sample (src/emitters/arealight.cpp:75-103, pointlight.cpp:47-63,
 * environmentmap.cpp:73-97) of emitters[emitter] on n (reference point, 2D sample) pairs — `in` holds 5
 * floats per record (ref.xyz, sample.xy). out gets 12 floats per record: lRec.wi, the returned
 * value (radiance / pdf), lRec.pdf, shadowRay.maxt, Emitter::eval for the sampled direction, 0. */
int  nb_emitter_probe(nb_ctx *ctx, uint32_t emitter, const float *in, uint64_t n, float *out);
/* Parity entry for the BSDF rows: BSDF::eval / pdf / sample (src/bsdf/{diffuse,mirror,dielectric,microfacet,
 * disney}.cpp) of bsdfs[bsdf] on n records of 10 floats { wi.xyz, wo.xyz, sample.xy, uv.xy } (local frame).
 * out gets 12 floats per record: eval(wi, wo) rgb for measure ESolidAngle, pdf(wi, wo), then for
 * sample(wi, sample): the returned weight rgb, the sampled wo.xyz, 1 if the measure is EDiscrete else 0,
 * and bsdf->pdf of the sampled direction. */
int  nb_bsdf_probe(nb_ctx *ctx, uint32_t bsdf, const float *in, uint64_t n, float *out);
/* Parity entry for the random streams: pcg32 (ext/pcg32/pcg32.h:51-66, 101-110) as the DEVICE runs it. out_u32 gets the
 * first n nextUInt() of a generator seeded with seed(initstate, initseq), out_f32 the first n nextFloat() of another
 * generator seeded the same way. (seed(42, 54) is "Round 1" of ext/pcg32/pcg32-demo.out.) */
int  nb_rng_probe(nb_ctx *ctx, uint64_t initstate, uint64_t initseq, uint32_t n, uint32_t *out_u32, float *out_f32);

/* RenderThread::renderThreadOptix loop (render.cpp:462-574). */
int  nb_render_begin(nb_ctx *ctx, const nb_render_params *params);  /* allocates + clears film */
/* OptixState::renderSubframe (OptixState.cpp:485-510): samples [first, first+count) of every
 * pixel; sample s of pixel p always uses the same random stream, whatever the launch split.
 * Only samples with (s % sample_stride) == sample_offset are rendered (multi-GPU sample slices;
 * 1,0 = all). Asynchronous (the default, persistent-kernel strategy): returns once the launch is queued on the
 * context's stream, so one host thread can keep several devices busy; nb_sync, nb_get_counters, nb_film_download*,
 * nb_film_clear and nb_render_begin wait for it, and a kernel failure is reported by whichever of them runs first.
 * (The optional wavefront strategy polls its queues and returns when the subframe is done. OptixState::renderSubframe
 * itself ends with cudaStreamSynchronize, OptixState.cpp:507 — call nb_sync for that behaviour.) */
int  nb_render_subframe(nb_ctx *ctx, uint32_t first, uint32_t count,
                        uint32_t sample_stride, uint32_t sample_offset);
/* One round of the outer sample loop (render.cpp:283-360) for a scene whose sampler is `adaptive`
 * (src/samplers/adaptive.cpp): for every block_size^2 block (NORI_BLOCK_SIZE_ADAPTIVE = 4, block.h:31) the loop
 * `while (sampler->computeVariance(block)) renderBlock(..)` of render.cpp:323-335 — initial_uniform (+1, the reference's
 * post-increment) passes of one sample per pixel, then passes whose samples are placed by the 3x3 luminance variance of the
 * block's previous pass (common.cpp:339-399) until the variance map stops settling — run as a feedback loop ON the device:
 * a decision kernel per pass (one warp per block) writes the list of samples, the render kernel consumes it and feeds the
 * blocks' own images back; the host reads one word per pass. All blocks advance together; a round ends when no block
 * goes on. Synchronous. The film accumulates like nb_render_subframe's (the two can be mixed); how many samples a round
 * draws depends on the image and comes back in `out`. block_size 2..8. round_cap: the sampler's limit of 1000 decisions per
 * block and round (adaptive.cpp:79), which only black blocks reach — they cost the reference 1000 passes per round, and
 * cost them here too unless the caller lowers the cap (0 = the reference's 1000). What is kept of the reference's quirks and
 * why is listed at the decision kernel in csrc/nb_render.cu. */
typedef struct nb_adaptive_info {
    uint64_t samples;         /* camera samples drawn in this round */
    uint64_t block_passes;    /* sum over blocks of the passes each one rendered */
    uint32_t passes;          /* device passes = the longest block loop */
    uint32_t blocks;          /* blocks in the image */
} nb_adaptive_info;
int  nb_render_adaptive_round(nb_ctx *ctx, uint32_t block_size, uint32_t initial_uniform, uint32_t round_cap, nb_adaptive_info *out);
/* Passes each block has rendered in adaptive rounds since they began (blocks row-major, ceil(width / block_size) per row) —
 * what AdaptiveSampler::writeVarianceMatrix's option (a) would show (adaptive.cpp:176-181); the tests compare its
 * distribution with the reference's. The counts restart when a round uses another block size. */
int  nb_adaptive_block_passes(nb_ctx *ctx, uint32_t *out, uint32_t n_blocks);
/* Host-only helper (no device needed): which samples a (first, count, stride, offset) slice holds:
 * *first_sample = smallest s >= first with s % stride == offset, *n_samples = how many such s lie
 * below first+count. The slices of offsets 0..stride-1 partition [first, first+count). */
int  nb_sample_slice(uint32_t first, uint32_t count, uint32_t stride, uint32_t offset,
                     uint32_t *first_sample, uint32_t *n_samples);
int  nb_sync(nb_ctx *ctx);
int  nb_film_clear(nb_ctx *ctx);
/* Film geometry: rows = H + 2b, cols = W + 2b, b = ceil(radius - 0.5) (block.cpp:57,73). */
int  nb_film_shape(nb_ctx *ctx, uint32_t *rows, uint32_t *cols, uint32_t *border);
/* Copies the accumulators into caller memory laid out like ImageBlock's storage (row-major
 * Color4f = {sum w*r, sum w*g, sum w*b, sum w}, block.h:49). normalize != 0: instead writes
 * H x W x 3 floats = ImageBlock::toBitmap() (block.cpp:76-82). */
int  nb_film_download(nb_ctx *ctx, float *out, int normalize);
/* The same for one layer (NB_LAYER_*); a layer that was not requested in nb_render_params.layers fails with NB_E_STATE. */
int  nb_film_download_layer(nb_ctx *ctx, int layer, float *out, int normalize);
/* Device address of the accumulators (rows*cols*4 floats), for callers that run their own collective. */
int  nb_film_device_ptr(nb_ctx *ctx, void **ptr, uint64_t *n_floats);
/* ---- multi-GPU: the film reduce (the film is linear, block.cpp:120-122) --------------
 * One process per GPU: rank 0 fills NB_COMM_ID_BYTES with nb_comm_unique_id, the launcher's own channel (MPI,
 * torch.distributed, a file) carries them to every rank, each rank calls nb_comm_init_rank on its context (collective:
 * returns when all n_ranks have joined). One process, n GPUs: nb_create_multi above.
 * nb_film_reduce: ncclReduce(sum, float, root = rank 0) of the film accumulators (and of the normal layer when the
 * render has one), in place, queued on the context's stream after the subframes already queued there — no host
 * synchronisation; a following nb_film_download on rank 0 returns the sum, and the other ranks' films are cleared (their
 * samples moved to the root, so render-more-then-reduce-again never counts a sample twice). With one rank it does nothing. In an
 * nb_create_multi group ONE call (on any member) reduces the whole group. NCCL is loaded at run time (libnccl.so.2);
 * without it these entries fail with NB_E_UNSUPPORTED and everything else still works. */
#define NB_COMM_ID_BYTES 128
int  nb_comm_unique_id(void *id);
int  nb_comm_init_rank(nb_ctx *ctx, int n_ranks, int rank, const void *id);
int  nb_comm_info(nb_ctx *ctx, int *rank, int *n_ranks, int *nccl_version);
int  nb_film_reduce(nb_ctx *ctx);
int  nb_get_counters(nb_ctx *ctx, nb_counters *out);
/* Device-side stopwatch on the context's own stream (the stream every kernel, copy and reduce of this library is
 * queued on — a caller's events on another stream do not see that work): nb_timer_mark(ctx, 0) ... queued work ...
 * nb_timer_mark(ctx, 1); nb_timer_elapsed waits for mark 1 and returns the CUDA-event time between the two. */
int  nb_timer_mark(nb_ctx *ctx, int which);
int  nb_timer_elapsed(nb_ctx *ctx, float *ms);
/* options: "profile" (per-kernel event timers of the wavefront pipeline), "sort" (reserved),
 * "strategy" (NB_STRATEGY_*: how path_mis / path_mats are executed; the direct_* integrators
 * always run as one K_tile launch), "flat_scan" (-1 auto, 0 always walk the tree, 1 always scan
 * the primitive list: same hits, see nb_trace.cuh), "walk" (1: tree scenes render path_mis / path_mats with the
 * per-lane state machine K_tile_walk), "ordered" (1: nb_trace_rays enters the child nearer to the ray first and breaks
 * exact-t ties by leaf-order slot, nb_traverse_ordered; 0, default: the reference's left-first order with skip links —
 * both return the reference's bits), "wide" (1, default: the render kernels traverse tree scenes through the PRODUCTION
 * tree — the binary tree collapsed to 4-wide nodes, children entered near-first, subtrees behind the current hit
 * culled; every primitive test and the exact-t tie rule are the reference's, so images are the same within Monte-Carlo
 * noise and hits agree except where a box test rounds differently, measured >= 99.9999 % (tests/test_wide_gpu.py);
 * 0: parity mode, the reference's own visiting order everywhere), "trace_wide" (1: nb_trace_rays answers through the
 * production tree too, to measure that hit parity; default 0 = bit-exact reference order), "trace_fast" (1: nb_trace_rays
 * tests triangles with fused multiply-adds — an experiment kept to MEASURE what contraction costs in parity: 3 of 24 656
 * primitive ids differ on the table scene and rays through shared edges can slip between triangles, which fails the
 * reference's own t-tests, so the render kernels never use it; default 0) */
enum { NB_STRATEGY_AUTO = 0, NB_STRATEGY_WAVEFRONT = 1, NB_STRATEGY_TILE = 2 };
int  nb_set_option(nb_ctx *ctx, const char *name, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* NORIB200_H */
