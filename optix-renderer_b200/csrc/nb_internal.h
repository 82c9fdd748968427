/* nb_internal.h — private to libnorib200.so: context, device scene, error plumbing. */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/norib200.h"

/* ---- device-side layouts ------------------------------------------------------------------
 * Node (32 B, two 16-byte loads): the reference's tree in the reference's order, re-encoded for
 * stackless traversal. The reference walks left-first with an explicit stack (bvh.cpp:428-451);
 * in its pre-order array "pop the stack" always lands on one fixed node per subtree, so we store
 * that node as `skip` and need no stack at all. Visiting order is unchanged.
 *   A = { skip, leaf_start (0xffffffff = inner), bmin.x, bmin.y }   B = { bmin.z, bmax.xyz }
 * Primitive slot (48 B, three 16-byte loads), in leaf order (BVH::m_indices order):
 *   triangle: { p0.xyz, e1.x } { e1.yz, e2.xy } { e2.z, bits(tag), bits(local prim), 0 }
 *   sphere:   { c.xyz, r }     { 0 }            { 0,    bits(tag), 0, 0 }
 *   tag = shape id | NB_SLOT_SPHERE | NB_SLOT_LAST (last slot of its leaf)
 * e1 = p1 - p0 and e2 = p2 - p0 are the same fp32 subtractions Mesh::rayIntersect does per test
 * (mesh.cpp:107), done once. */
#define NB_SLOT_SPHERE 0x80000000u
#define NB_SLOT_LAST   0x40000000u
#define NB_SLOT_SHAPE_MASK 0x3fffffffu
#define NB_NODE_END 0xffffffffu
/* A.y of an inner node: bit 31 set; bits 29-30 = the axis along which its children's box centres are
 * furthest apart; bit 28 = the LEFT child is the upper one along that axis; bits 0-27 = index of the
 * right child (the left child is idx + 1). Only the ordered traversal reads the low bits. */
#define NB_NODE_INNER_BIT 0x80000000u
#define NB_NODE_RIGHT_MASK 0x0fffffffu
#define NB_NODE_EMPTY_LEAF 0x7fffffffu
#define NB_STACK_DEPTH 64u              /* as the reference's traversal stack (bvh.cpp:413) */
__host__ __device__ inline uint32_t nb_inner_word(uint32_t right, const float lmin[3], const float lmax[3], const float rmin[3], const float rmax[3]) {
    int axis = 0; float best = -1.f; bool left_upper = false;
    for (int a = 0; a < 3; ++a) {
        const float dl = (rmin[a] + rmax[a]) - (lmin[a] + lmax[a]);
        const float ad = dl < 0.f ? -dl : dl;
        if (ad > best) { best = ad; axis = a; left_upper = dl < 0.f; }
    }
    return NB_NODE_INNER_BIT | ((uint32_t) axis << 29) | (left_upper ? (1u << 28) : 0u) | (right & NB_NODE_RIGHT_MASK);
}
#define NB_FLAT_MAX_SLOTS 24u
/* ---- production tree (not the reference's layout) ----------------------------------
 * The binary tree collapsed to 4-wide nodes on the host (greedy by surface area). A wide node is 4 child records of
 * 32 B = 128 B; child record c at float4 [8 * node + 2 * c]:
 *   A = { bits(ref), 0, bmin.x, bmin.y }   B = { bmin.z, bmax.xyz }       (the binary node's own fp32 box)
 *   ref: NB_WIDE_EMPTY = no child; bit 31 set = leaf, low bits = its first primitive slot (the leaf ends at the slot
 *   tagged NB_SLOT_LAST, as in the binary tree — slots stay in the reference's m_indices order, which is what the
 *   exact-t tie rule needs); else the index of the child wide node.
 * Traversal is near-first with a short stack of (ref, entry distance), so the closest hit shrinks the ray early and
 * whole subtrees are culled when popped; per ray it touches ~3x fewer nodes than the reference's left-first order. */
#define NB_WIDE_LEAF 0x80000000u
#define NB_WIDE_EMPTY 0xffffffffu
#define NB_WIDE_STACK 64u               /* entries a ray may hold; trees deeper than (NB_WIDE_STACK - 1) / 3 levels keep the binary walk */

struct NbDeviceScene {
    const float4 *nodes = nullptr;      /* 2 per node */
    const float4 *slots = nullptr;      /* 3 per primitive slot */
    uint32_t n_nodes = 0, n_slots = 0;
    uint32_t flat = 0;                  /* queries scan the slots instead of walking the tree (nb_trace.cuh) */
    uint32_t ordered = 0;               /* nb_trace_rays walks near child first (nb_traverse_ordered) instead of left first */
    const float4 *wide = nullptr;       /* production tree: 4-wide nodes collapsed from the binary tree (8 float4 per node, see NB_WIDE_*) */
    uint32_t n_wide = 0;
    uint32_t use_wide = 0;              /* queries walk the wide tree near-first (nb_traverse_wide) — the production mode; 0 = the reference's order */
    uint32_t ext = 0;                   /* scene uses textures, normal maps, an environment map, a heterogeneous medium, the disney BSDF or a spot / directional light: the EXT kernel variants render it */
    /* shading data */
    const nb_shape *shapes = nullptr;
    const float *positions = nullptr, *normals = nullptr, *uvs = nullptr, *tangents = nullptr, *bitangents = nullptr;
    const uint32_t *faces = nullptr;
    const nb_bsdf *bsdfs = nullptr;
    const nb_emitter *emitters = nullptr;
    const float *emitter_cdf = nullptr, *area_cdf = nullptr;
    const nb_medium *media = nullptr;
    const float *density = nullptr;     /* voxels of heterogeneous media */
    const nb_texture *textures = nullptr;
    const float *texels = nullptr;      /* RGBA, 16-byte aligned */
    const int32_t *medium_emitter = nullptr;   /* per medium: index of its volume light (Medium::getEmitter), -1 = none; null when no medium has one */
    const float *params = nullptr;      /* nb_scene_desc.params: disney / spot / directional parameters */
    uint32_t n_shapes = 0, n_bsdfs = 0, n_emitters = 0, n_media = 0;
    int32_t envmap_emitter = -1, ambient_medium = -1;
    nb_camera camera;
#ifdef NB_PARAM_SLOTS
    /* Experiment kept behind a macro, not adopted: the primitive slots of a flat scene (<= NB_FLAT_MAX_SLOTS) once more
     * inside the kernel parameters, so that the flat scan's warp-uniform slot reads become constant-bank broadcasts (LDCU
     * into uniform registers) instead of three per-lane LDG.128. B200, M samples/s, global loads -> parameter space:
     * cbox path_mis 2019 -> 1817, cbox_val_mis 988 -> 847, path_mats 2756 -> 2691, homog_aniso 3005 -> 2926. */
    float4 flat_slots[3 * NB_FLAT_MAX_SLOTS];
#endif
    uint32_t flat_in_params = 0;
    uint32_t fast_hits = 0;             /* nb_trace_rays only ("trace_fast"): the production triangle test instead of the reference's */
};

struct NbRenderState;   /* nb_render.cu */

struct nb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string error;
    /* host copies kept for nb_build_bvh / hit translation */
    std::vector<nb_shape> h_shapes;
    std::vector<float> h_positions;
    std::vector<uint32_t> h_faces;
    std::vector<uint32_t> h_shape_prim_offset;   /* BVH::m_shapeOffset */
    std::vector<uint32_t> h_slot_tag, h_slot_prim; /* per leaf-order slot: shape tag, local prim */
    bool scene_ready = false, bvh_ready = false, render_begun = false;
    NbDeviceScene dev;
    std::vector<void *> scene_allocs, bvh_allocs;
    NbRenderState *render = nullptr;
    int sm_count = 148;
    double bvh_build_ms = 0.0;          /* device time of the last NB_BVH_DEVICE build */
    int opt_profile = 0, opt_sort = 0, opt_strategy = 0;
    int opt_walk = 1;                   /* tree scenes: K_tile_walk (per-lane state machine) instead of lock-step K_tile */
    int opt_ordered = 0;                /* 1: nb_trace_rays uses the near-child-first walk */
    int opt_trace_fast = 0;             /* 1: nb_trace_rays uses the production triangle test too (to measure its parity) */
    int opt_trace_wide = 0;             /* 1: nb_trace_rays answers through the wide tree too (to measure its hit parity) */
    int opt_wide = 1;                   /* 1: tree scenes are traversed through the wide tree (production); 0: the reference's order (parity mode) */
    uint32_t wide_depth = 0;            /* levels of the wide tree */
    int opt_flat = -1;                  /* -1: flat scan iff the scene has at most NB_FLAT_MAX_SLOTS primitives */
    bool flat_ok = true;                /* leaves of the imported tree cover the slot list in pre-order: the flat scan visits what the walk visits */
    /* multi-GPU film reduce (nb_comm.cu): one NCCL communicator per context */
    void *comm = nullptr;               /* ncclComm_t */
    int comm_rank = 0, comm_size = 1;
    struct NbGroup *group = nullptr;    /* contexts created together by nb_create_multi (one process, n devices) */
    cudaEvent_t timer[2] = { nullptr, nullptr };   /* nb_timer_mark / nb_timer_elapsed */
};

int nb_fail(nb_ctx *ctx, int code, const char *fmt, ...);
inline bool nb_scene_flat(const nb_ctx *ctx) {
    if (!ctx->flat_ok) return false;
    if (ctx->opt_flat >= 0) return ctx->opt_flat != 0;
    return ctx->dev.n_slots <= NB_FLAT_MAX_SLOTS;
}
#define NB_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return nb_fail(ctx, NB_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

/* nb_trace.cu */
void nb_launch_trace_hits(const NbDeviceScene &scene, const float4 *rays, uint64_t n, int any_hit,
                          float4 *out, uint32_t *out_any, int sm_count, cudaStream_t stream);

/* nb_render.cu */
int nb_render_free(nb_ctx *ctx);
/* waits for the subframes queued by nb_render_subframe and folds their timers / counters into the totals */
int nb_render_finish(nb_ctx *ctx);
/* film accumulators of the current render: data (+ normal layer or null), float count of one layer */
int nb_render_film(nb_ctx *ctx, float **data, float **normal, uint64_t *n_floats);

/* nb_comm.cu */
void nb_comm_free(nb_ctx *ctx);

/* nb_bvh.cu */
void nb_collapse_wide(const nb_bvh_node *nodes, uint32_t n, std::vector<float4> &wide, uint32_t &depth);
void nb_update_wide(nb_ctx *ctx);

/* nb_bvh_build.cu */
int nb_build_bvh_device(nb_ctx *ctx, int sah);

/* nb_mega.cu */
struct NbFilm;
struct NbPlan;
cudaError_t nb_launch_tile(const NbDeviceScene &sc, const NbFilm &film, const NbPlan &plan, uint32_t height, uint32_t passes,
                           uint32_t integrator, uint32_t *item_counter, unsigned long long *stats, int sm_count,
                           cudaStream_t stream, bool walk, const uint32_t *list = nullptr, const uint32_t *list_ctl = nullptr);
