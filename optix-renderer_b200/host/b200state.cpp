#include "b200state.h"

namespace nori {

B200State::B200State(int device) {
    int rc = nb_create(device, &m_ctx);
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(nullptr));
}

B200State::~B200State() { nb_destroy(m_ctx); }

void B200State::check(int rc) const {
    if (rc != NB_OK) throw NoriException("b200: %s", nb_last_error(m_ctx));
}

bool B200State::preRender(const Scene &scene, uint64_t seed, uint32_t poolPaths, uint32_t layers) {
    if (!scene.flatten(m_flat, &m_why)) return false;
    nb_scene_desc d = m_flat.desc();
    int rc = nb_upload_scene(m_ctx, &d);
    if (rc == NB_E_UNSUPPORTED) { m_why = nb_last_error(m_ctx); return false; }
    check(rc);
    nb_bvh_import imp = { (uint32_t) m_flat.bvh_nodes.size(), m_flat.bvh_nodes.data(),
                          (uint32_t) m_flat.bvh_indices.size(), m_flat.bvh_indices.data() };
    check(nb_build_bvh(m_ctx, NB_BVH_IMPORT, &imp));
    nb_render_params p = m_flat.render;
    p.seed = seed; p.pool_paths = poolPaths; p.layers = layers;
    rc = nb_render_begin(m_ctx, &p);
    if (rc == NB_E_UNSUPPORTED) { m_why = nb_last_error(m_ctx); return false; }
    check(rc);
    return true;
}

void B200State::renderSubframe(uint32_t currentSample, uint32_t samplesPerLaunch, uint32_t stride, uint32_t offset) {
    check(nb_render_subframe(m_ctx, currentSample, samplesPerLaunch, stride, offset));
}

void B200State::download(std::vector<float> &out, bool normalize, uint32_t *rows, uint32_t *cols, int layer) {
    uint32_t r = 0, c = 0, b = 0;
    check(nb_film_shape(m_ctx, &r, &c, &b));
    if (normalize) { r -= 2 * b; c -= 2 * b; }
    out.resize((size_t) r * c * (normalize ? 3 : 4));
    check(nb_film_download_layer(m_ctx, layer, out.data(), normalize ? 1 : 0));
    if (rows) *rows = r;
    if (cols) *cols = c;
}

void B200State::clear() { check(nb_film_clear(m_ctx)); }

nb_counters B200State::counters() const {
    nb_counters c;
    check(nb_get_counters(m_ctx, &c));
    return c;
}

} // namespace nori
