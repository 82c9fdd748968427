/* nb_bvh.cu — nb_build_bvh (import of the caller's BVH) and the parity trace entries. */
#include "nb_internal.h"
#include <algorithm>
#include <cstring>

/* Collapses the binary tree (reference node layout) into 4-wide nodes: a wide node starts from a binary node's two
 * children and, while it has room, replaces the inner child with the largest surface area by that child's two
 * children. Child boxes are the binary nodes' own boxes; leaves keep their slot ranges. Wide nodes are numbered in
 * depth-first order (a node's subtree is contiguous). Returns the number of levels. */
void nb_collapse_wide(const nb_bvh_node *nodes, uint32_t n, std::vector<float4> &wide, uint32_t &depth) {
    wide.clear(); depth = 0;
    if (n == 0) return;
    auto area = [&](uint32_t i) {
        const nb_bvh_node &b = nodes[i];
        const float dx = b.bmax[0] - b.bmin[0], dy = b.bmax[1] - b.bmin[1], dz = b.bmax[2] - b.bmin[2];
        return dx * dy + dy * dz + dz * dx;
    };
    auto is_leaf = [&](uint32_t i) { return (nodes[i].word0 & 1u) != 0; };
    struct Work { uint32_t bin, wide, level; };
    std::vector<Work> todo;
    auto new_node = [&]() {
        const uint32_t w = (uint32_t) (wide.size() / 8);
        wide.resize(wide.size() + 8, make_float4(0.f, 0.f, 0.f, 0.f));
        for (int c = 0; c < 4; ++c) { const uint32_t e = NB_WIDE_EMPTY; memcpy(&wide[(size_t) w * 8 + 2 * c].x, &e, 4); }
        return w;
    };
    auto put_child = [&](uint32_t w, int c, uint32_t bin, uint32_t ref) {
        const nb_bvh_node &b = nodes[bin];
        float4 A = make_float4(0.f, 0.f, b.bmin[0], b.bmin[1]);
        memcpy(&A.x, &ref, 4);
        wide[(size_t) w * 8 + 2 * c] = A;
        wide[(size_t) w * 8 + 2 * c + 1] = make_float4(b.bmin[2], b.bmax[0], b.bmax[1], b.bmax[2]);
    };
    const uint32_t root = new_node();
    if (is_leaf(0)) {                               /* a one-leaf tree: a wide root with that leaf as its only child */
        if ((nodes[0].word0 >> 1) > 0) put_child(root, 0, 0, NB_WIDE_LEAF | nodes[0].word1);
        depth = 1;
        return;
    }
    todo.push_back({ 0u, root, 1u });
    while (!todo.empty()) {
        const Work wk = todo.back();
        todo.pop_back();
        depth = std::max(depth, wk.level);
        uint32_t kids[4] = { wk.bin + 1, nodes[wk.bin].word1, 0, 0 };
        int nk = 2;
        while (nk < 4) {
            int best = -1; float best_area = -1.f;
            for (int k = 0; k < nk; ++k)
                if (!is_leaf(kids[k]) && area(kids[k]) > best_area) { best = k; best_area = area(kids[k]); }
            if (best < 0) break;
            const uint32_t b = kids[best];
            kids[best] = b + 1;
            kids[nk++] = nodes[b].word1;
        }
        /* children that are inner nodes get their wide node now, in reverse so that the first is processed first */
        uint32_t refs[4];
        for (int k = 0; k < nk; ++k) {
            if (is_leaf(kids[k])) refs[k] = (nodes[kids[k]].word0 >> 1) ? (NB_WIDE_LEAF | nodes[kids[k]].word1) : NB_WIDE_EMPTY;
            else refs[k] = new_node();
        }
        for (int k = nk - 1; k >= 0; --k)
            if (!is_leaf(kids[k])) todo.push_back({ kids[k], refs[k], wk.level + 1 });
        for (int k = 0; k < nk; ++k)
            if (refs[k] != NB_WIDE_EMPTY) put_child(wk.wide, k, kids[k], refs[k]);
    }
}

/* which traversal the render kernels use for this tree */
void nb_update_wide(nb_ctx *ctx) {
    const bool fits = ctx->dev.wide && ctx->dev.n_wide && 3u * ctx->wide_depth + 1u <= NB_WIDE_STACK;
    ctx->dev.use_wide = (ctx->opt_wide && fits) ? 1u : 0u;
}

extern "C" {

/* Re-encodes BVH::m_nodes / m_indices (include/nori/bvh.h:127-164) into the stackless device
 * layout of nb_internal.h. The tree and the order in which nodes and primitives are visited
 * are unchanged; only "pop the stack" is precomputed as a skip link. */
int nb_build_bvh(nb_ctx *ctx, int mode, const nb_bvh_import *imp) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->scene_ready) return nb_fail(ctx, NB_E_STATE, "nb_build_bvh: upload a scene first");
    if (mode == NB_BVH_DEVICE || mode == NB_BVH_DEVICE_SAH) {
        NB_CUDA(ctx, cudaSetDevice(ctx->device));
        return nb_build_bvh_device(ctx, mode == NB_BVH_DEVICE_SAH ? 1 : 0);
    }
    if (mode != NB_BVH_IMPORT) return nb_fail(ctx, NB_E_UNSUPPORTED, "nb_build_bvh: mode %d not available", mode);
    if (!imp || !imp->nodes || !imp->indices || imp->n_nodes == 0)
        return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: empty import");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint32_t n = imp->n_nodes, ni = imp->n_indices;
    const uint32_t total_prims = ctx->h_shape_prim_offset.back();
    if (n > NB_NODE_RIGHT_MASK) return nb_fail(ctx, NB_E_UNSUPPORTED, "nb_build_bvh: %u nodes exceed the node index range", n);
    if (ni != total_prims) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: %u indices for %u primitives", ni, total_prims);

    std::vector<uint32_t> skip(n, NB_NODE_END);
    std::vector<uint8_t> seen(n, 0);
    std::vector<uint32_t> stack;
    stack.push_back(0);
    uint32_t leaf_end = 0;          /* end of the last leaf range met in pre-order */
    bool leaves_in_order = true;
    while (!stack.empty()) {
        const uint32_t i = stack.back();
        stack.pop_back();
        if (i >= n || seen[i]) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: node %u out of range or visited twice", i);
        seen[i] = 1;
        const nb_bvh_node &nd = imp->nodes[i];
        if (nd.word0 & 1u) {
            const uint32_t size = nd.word0 >> 1, start = nd.word1;
            if ((uint64_t) start + size > ni) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: leaf %u range outside indices", i);
            if (size) { if (start < leaf_end) leaves_in_order = false; leaf_end = start + size; }
        } else {
            const uint32_t left = i + 1, right = nd.word1;
            if (left >= n || right >= n) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: inner node %u has a child out of range", i);
            skip[left] = right;
            skip[right] = skip[i];
            stack.push_back(right);
            stack.push_back(left);
        }
    }

    /* primitive slots in leaf (m_indices) order */
    std::vector<float4> slots((size_t) ni * 3);
    for (uint32_t k = 0; k < ni; ++k) {
        const uint32_t gid = imp->indices[k];
        if (gid >= total_prims) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: primitive id %u out of range", gid);
        /* BVH::findShape, bvh.h:105-109 */
        const uint32_t sid = (uint32_t) (std::upper_bound(ctx->h_shape_prim_offset.begin(), ctx->h_shape_prim_offset.end(), gid)
                                         - ctx->h_shape_prim_offset.begin()) - 1;
        const nb_shape &sh = ctx->h_shapes[sid];
        const uint32_t local = gid - ctx->h_shape_prim_offset[sid];
        float4 a = make_float4(0, 0, 0, 0), b = a, c = a;
        uint32_t tag = sid;
        if (sh.type == NB_SHAPE_SPHERE) {
            a = make_float4(sh.center[0], sh.center[1], sh.center[2], sh.radius);
            tag |= NB_SLOT_SPHERE;
        } else {
            const uint32_t *f = &ctx->h_faces[((size_t) sh.face_offset + local) * 3];
            const float *p0 = &ctx->h_positions[((size_t) sh.vertex_offset + f[0]) * 3];
            const float *p1 = &ctx->h_positions[((size_t) sh.vertex_offset + f[1]) * 3];
            const float *p2 = &ctx->h_positions[((size_t) sh.vertex_offset + f[2]) * 3];
            /* edge1 = p1 - p0, edge2 = p2 - p0 (mesh.cpp:107) */
            a = make_float4(p0[0], p0[1], p0[2], p1[0] - p0[0]);
            b = make_float4(p1[1] - p0[1], p1[2] - p0[2], p2[0] - p0[0], p2[1] - p0[1]);
            c.x = p2[2] - p0[2];
        }
        memcpy(&c.y, &tag, 4);
        memcpy(&c.z, &local, 4);
        slots[(size_t) k * 3] = a; slots[(size_t) k * 3 + 1] = b; slots[(size_t) k * 3 + 2] = c;
    }

    std::vector<float4> nodes((size_t) n * 2);
    for (uint32_t i = 0; i < n; ++i) {
        const nb_bvh_node &nd = imp->nodes[i];
        uint32_t w1 = 0;
        if (!(nd.word0 & 1u)) {
            const nb_bvh_node &l = imp->nodes[i + 1], &r = imp->nodes[nd.word1];
            w1 = nb_inner_word(nd.word1, l.bmin, l.bmax, r.bmin, r.bmax);
        } else {
            const uint32_t size = nd.word0 >> 1, start = nd.word1;
            if (size == 0) w1 = NB_NODE_EMPTY_LEAF;
            else {
                w1 = start;
                uint32_t tag;
                memcpy(&tag, &slots[(size_t) (start + size - 1) * 3 + 2].y, 4);
                tag |= NB_SLOT_LAST;
                memcpy(&slots[(size_t) (start + size - 1) * 3 + 2].y, &tag, 4);
            }
        }
        float4 A, B;
        memcpy(&A.x, &skip[i], 4);
        memcpy(&A.y, &w1, 4);
        A.z = nd.bmin[0]; A.w = nd.bmin[1];
        B = make_float4(nd.bmin[2], nd.bmax[0], nd.bmax[1], nd.bmax[2]);
        nodes[(size_t) i * 2] = A; nodes[(size_t) i * 2 + 1] = B;
    }

    /* production tree: the same binary tree collapsed to 4-wide nodes (nb_internal.h, NB_WIDE_*) */
    std::vector<float4> wide;
    uint32_t wide_depth = 0;
    nb_collapse_wide(imp->nodes, n, wide, wide_depth);

    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }     /* queued subframes still walk the old tree */
    ctx->bvh_ready = false;                      /* a failure below must not leave a context pointing at freed memory */
    ctx->dev.nodes = nullptr; ctx->dev.slots = nullptr; ctx->dev.n_nodes = 0; ctx->dev.n_slots = 0;
    ctx->dev.wide = nullptr; ctx->dev.n_wide = 0; ctx->dev.use_wide = 0;
    for (void *p : ctx->bvh_allocs) cudaFree(p);
    ctx->bvh_allocs.clear();
    void *dn = nullptr, *ds = nullptr;
    NB_CUDA(ctx, cudaMalloc(&dn, nodes.size() * sizeof(float4)));
    ctx->bvh_allocs.push_back(dn);
    NB_CUDA(ctx, cudaMalloc(&ds, std::max<size_t>(slots.size(), 1) * sizeof(float4)));
    ctx->bvh_allocs.push_back(ds);
    void *dw = nullptr;
    NB_CUDA(ctx, cudaMalloc(&dw, std::max<size_t>(wide.size(), 1) * sizeof(float4)));
    ctx->bvh_allocs.push_back(dw);
    NB_CUDA(ctx, cudaMemcpyAsync(dw, wide.data(), wide.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    NB_CUDA(ctx, cudaMemcpyAsync(dn, nodes.data(), nodes.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    NB_CUDA(ctx, cudaMemcpyAsync(ds, slots.data(), slots.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->dev.nodes = (const float4 *) dn; ctx->dev.slots = (const float4 *) ds;
    ctx->dev.n_nodes = n; ctx->dev.n_slots = ni;
    ctx->dev.wide = (const float4 *) dw; ctx->dev.n_wide = (uint32_t) (wide.size() / 8);
    ctx->wide_depth = wide_depth;
    nb_update_wide(ctx);
    /* the flat scan tests slots 0..n-1 in order; that is the walk's order only when the caller's leaves come in
     * increasing index order (true of BVH::build's output; a foreign tree that breaks it is walked instead) */
    ctx->flat_ok = leaves_in_order;
    ctx->dev.flat = nb_scene_flat(ctx) ? 1u : 0u;
    ctx->dev.flat_in_params = 0;
#ifdef NB_PARAM_SLOTS
    if (ni <= NB_FLAT_MAX_SLOTS) {
        memcpy(ctx->dev.flat_slots, slots.data(), slots.size() * sizeof(float4));
        ctx->dev.flat_in_params = 1;
    }
#endif
    ctx->h_slot_tag.resize(ni); ctx->h_slot_prim.resize(ni);
    for (uint32_t k = 0; k < ni; ++k) {
        memcpy(&ctx->h_slot_tag[k], &slots[(size_t) k * 3 + 2].y, 4);
        memcpy(&ctx->h_slot_prim[k], &slots[(size_t) k * 3 + 2].z, 4);
    }
    ctx->bvh_build_ms = 0.0;
    ctx->bvh_ready = true;
    return NB_OK;
}

int nb_trace_rays(nb_ctx *ctx, const nb_ray *rays, uint64_t n, int any_hit, nb_hit *out) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->bvh_ready) return nb_fail(ctx, NB_E_STATE, "nb_trace_rays: no BVH (call nb_build_bvh)");
    if (n == 0) return NB_OK;
    if (!rays || !out) return nb_fail(ctx, NB_E_INVALID, "nb_trace_rays: NULL buffer");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    std::vector<float4> packed((size_t) n * 2);
    for (uint64_t i = 0; i < n; ++i) {
        const nb_ray &r = rays[i];
        packed[2 * i] = make_float4(r.o[0], r.o[1], r.o[2], r.mint);
        packed[2 * i + 1] = make_float4(r.d[0], r.d[1], r.d[2], r.maxt);
    }
    float4 *d_rays = nullptr, *d_out = nullptr;
    uint32_t *d_any = nullptr;
    int rc = NB_OK;
    std::vector<float4> h_out((size_t) n);
    std::vector<uint32_t> h_any(any_hit ? (size_t) n : 0);
    cudaError_t e;
    if ((e = cudaMalloc((void **) &d_rays, packed.size() * sizeof(float4))) != cudaSuccess ||
        (e = cudaMalloc((void **) &d_out, (size_t) n * sizeof(float4))) != cudaSuccess ||
        (any_hit && (e = cudaMalloc((void **) &d_any, (size_t) n * 4)) != cudaSuccess)) {
        rc = nb_fail(ctx, NB_E_NOMEM, "nb_trace_rays: %s", cudaGetErrorString(e));
    } else {
        cudaMemcpyAsync(d_rays, packed.data(), packed.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream);
        NbDeviceScene sc = ctx->dev;               /* the parity entry walks in the reference's order unless "trace_wide" asks for the production tree */
        sc.use_wide = (ctx->opt_trace_wide && sc.wide) ? 1u : 0u;
        sc.fast_hits = ctx->opt_trace_fast ? 1u : 0u;
        nb_launch_trace_hits(sc, d_rays, n, any_hit, d_out, d_any, ctx->sm_count, ctx->stream);
        cudaMemcpyAsync(h_out.data(), d_out, (size_t) n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream);
        if (any_hit) cudaMemcpyAsync(h_any.data(), d_any, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream);
        e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) rc = nb_fail(ctx, NB_E_CUDA, "nb_trace_rays: %s", cudaGetErrorString(e));
    }
    cudaFree(d_rays); cudaFree(d_out); cudaFree(d_any);
    if (rc != NB_OK) return rc;
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t slot;
        memcpy(&slot, &h_out[i].w, 4);
        nb_hit &h = out[i];
        h.t = h_out[i].x; h.u = h_out[i].y; h.v = h_out[i].z;
        if (slot == 0xffffffffu) { h.shape = -1; h.prim = -1; }
        else { h.shape = (int32_t) (ctx->h_slot_tag[slot] & NB_SLOT_SHAPE_MASK); h.prim = (int32_t) ctx->h_slot_prim[slot]; }
        h.any_hit = any_hit ? h_any[i] : 0u;
    }
    return NB_OK;
}

int nb_trace_rays_device(nb_ctx *ctx, const void *rays_device, uint64_t n, int any_hit, void *out_device,
                         int repeats, float *ms_per_launch) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->bvh_ready) return nb_fail(ctx, NB_E_STATE, "nb_trace_rays_device: no BVH");
    if (!rays_device || !out_device || repeats < 1) return nb_fail(ctx, NB_E_INVALID, "nb_trace_rays_device: bad argument");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *d_any = nullptr;
    if (any_hit) NB_CUDA(ctx, cudaMalloc((void **) &d_any, (size_t) n * 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, ctx->stream);
    NbDeviceScene sc = ctx->dev;
    sc.use_wide = (ctx->opt_trace_wide && sc.wide) ? 1u : 0u;
    sc.fast_hits = ctx->opt_trace_fast ? 1u : 0u;
    for (int r = 0; r < repeats; ++r)
        nb_launch_trace_hits(sc, (const float4 *) rays_device, n, any_hit, (float4 *) out_device, d_any, ctx->sm_count, ctx->stream);
    cudaEventRecord(e1, ctx->stream);
    cudaError_t e = cudaEventSynchronize(e1);
    if (e == cudaSuccess) e = cudaGetLastError();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d_any);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_trace_rays_device: %s", cudaGetErrorString(e));
    if (ms_per_launch) *ms_per_launch = ms / repeats;
    return NB_OK;
}

int nb_bvh_stats(nb_ctx *ctx, uint32_t *n_nodes, uint32_t *n_slots, double *build_ms) {
    if (!ctx || !ctx->bvh_ready) return nb_fail(ctx, NB_E_STATE, "nb_bvh_stats: no BVH");
    if (n_nodes) *n_nodes = ctx->dev.n_nodes;
    if (n_slots) *n_slots = ctx->dev.n_slots;
    if (build_ms) *build_ms = ctx->bvh_build_ms;
    return NB_OK;
}

} /* extern "C" */
