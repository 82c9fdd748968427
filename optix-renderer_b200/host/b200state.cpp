#include "b200state.h"
#include <cstdlib>

namespace nori {

B200State::B200State(int device) {
    int rc = nb_create(device, &m_ctx);
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(nullptr));
    m_all.push_back(m_ctx);
}

B200State::B200State(const std::vector<int> &devices) {
    if (devices.empty()) throw NoriException("b200: empty device list");
    m_all.assign(devices.size(), nullptr);
    int rc = nb_create_multi(devices.data(), (int) devices.size(), m_all.data());
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(nullptr));
    m_ctx = m_all[0];
}

B200State::~B200State() { for (nb_ctx *c : m_all) nb_destroy(c); }

void B200State::check(int rc) const {
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(m_ctx));
}

static void checkOn(nb_ctx *c, int rc) {
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(c));
}

bool B200State::preRender(const Scene &scene, uint64_t seed, uint32_t poolPaths, uint32_t layers) {
    int mode = m_treeMode;
    if (const char *env = getenv("NB_BVH_BUILDER")) {
        const std::string b(env);
        if (b == "host") mode = NB_BVH_IMPORT; else if (b == "lbvh") mode = NB_BVH_DEVICE; else if (b == "sah") mode = NB_BVH_DEVICE_SAH;
        else throw NoriException("NB_BVH_BUILDER=%s: host, lbvh or sah", b);
    }
    if (!scene.flatten(m_flat, &m_why, mode == NB_BVH_IMPORT)) return false;
    nb_scene_desc d = m_flat.desc();
    nb_bvh_import imp = { (uint32_t) m_flat.bvh_nodes.size(), m_flat.bvh_nodes.data(),
                          (uint32_t) m_flat.bvh_indices.size(), m_flat.bvh_indices.data() };
    nb_render_params p = m_flat.render;
    p.seed = seed; p.pool_paths = poolPaths; p.layers = layers;
    if (m_boxFilter) {
        p.filter_radius = 0.5f;
        for (int i = 0; i < NB_FILTER_RESOLUTION; ++i) p.filter_table[i] = 1.f;
        p.filter_table[NB_FILTER_RESOLUTION] = 0.f;
    }
    for (nb_ctx *c : m_all) {                        /* the scene is replicated per GPU */
        int rc = nb_upload_scene(c, &d);
        if (rc == NB_E_UNSUPPORTED) { m_why = nb_last_error(c); return false; }
        checkOn(c, rc);
        checkOn(c, nb_build_bvh(c, mode, mode == NB_BVH_IMPORT ? &imp : nullptr));
        rc = nb_render_begin(c, &p);
        if (rc == NB_E_UNSUPPORTED) { m_why = nb_last_error(c); return false; }
        checkOn(c, rc);
    }
    return true;
}

void B200State::renderSubframe(uint32_t currentSample, uint32_t samplesPerLaunch, uint32_t stride, uint32_t offset) {
    /* device g of n takes the samples s of the caller's slice with (s / stride) % n == g: slice (stride * n,
     * offset + g * stride). nb_render_subframe only queues the launch, so this loop keeps every device busy;
     * like OptixState::renderSubframe (OptixState.cpp:507) the call returns when the subframe is done. */
    const uint32_t n = (uint32_t) m_all.size();
    for (uint32_t g = 0; g < n; ++g)
        checkOn(m_all[g], nb_render_subframe(m_all[g], currentSample, samplesPerLaunch, stride * n, offset + g * stride));
    for (nb_ctx *c : m_all) checkOn(c, nb_sync(c));
}

nb_adaptive_info B200State::renderAdaptiveRound(uint32_t blockSize, uint32_t initialUniform, uint32_t roundCap) {
    if (m_all.size() > 1) throw NoriException("B200State: the adaptive sampler runs on one device (its feedback reads a block's own samples)");
    nb_adaptive_info info;
    check(nb_render_adaptive_round(m_ctx, blockSize, initialUniform, roundCap, &info));
    return info;
}

void B200State::download(std::vector<float> &out, bool normalize, uint32_t *rows, uint32_t *cols, int layer) {
    uint32_t r = 0, c = 0, b = 0;
    check(nb_film_shape(m_ctx, &r, &c, &b));
    if (normalize) { r -= 2 * b; c -= 2 * b; }
    out.resize((size_t) r * c * (normalize ? 3 : 4));
    if (m_all.size() > 1) check(nb_film_reduce(m_ctx));      /* one NCCL sum of the accumulators onto the first device */
    check(nb_film_download_layer(m_ctx, layer, out.data(), normalize ? 1 : 0));
    if (rows) *rows = r;
    if (cols) *cols = c;
}

void B200State::clear() { for (nb_ctx *c : m_all) checkOn(c, nb_film_clear(c)); }

nb_counters B200State::counters() const {
    nb_counters total;
    check(nb_get_counters(m_ctx, &total));
    for (size_t g = 1; g < m_all.size(); ++g) {
        nb_counters c;
        checkOn(m_all[g], nb_get_counters(m_all[g], &c));
        total.samples += c.samples; total.rays_closest += c.rays_closest; total.rays_any += c.rays_any;
        total.kernel_launches += c.kernel_launches; total.iterations += c.iterations; total.invalid_samples += c.invalid_samples;
        if (c.device_ms > total.device_ms) total.device_ms = c.device_ms;       /* devices run side by side */
    }
    return total;
}

} // namespace nori
