/* scene.cpp — Scene container (src/utils/scene.cpp:204-266 addChild, :59-151 defaults) and the
 * flatten step that OptixState::preRender performs in the reference
 * (include/nori/optix/OptixState.render.cpp:19-72): walk the plugins, call each flatten hook,
 * build the BVH, tabulate the reconstruction filter as ImageBlock::init does (block.cpp:58-64). */
#include "plugins.h"
#include "../../include/nb_scene_file.h"
#include <cstring>

namespace nori {

Scene::Scene(const PropertyList &) {}

Scene::~Scene() {
    for (Emitter *e : m_emitters) {                               /* shape- and medium-owned emitters die with their owner */
        bool owned = e->hasShape();
        for (Shape *s : m_shapes) owned |= s->getMediumEmitter() == e;
        if (!owned) delete e;
    }
    for (Shape *s : m_shapes) delete s;
    delete m_camera; delete m_integrator; delete m_sampler; delete m_ambientMedium;
    for (NoriObject *o : m_other) delete o;
}

void Scene::addChild(NoriObject *obj) {
    switch (obj->getClassType()) {
        case EShape: {
            Shape *shape = static_cast<Shape *>(obj);
            m_shapes.push_back(shape);
            if (shape->isEmitter()) m_emitters.push_back(shape->getEmitter());   /* scene.cpp:211-215 */
            if (shape->getMediumEmitter()) m_emitters.push_back(shape->getMediumEmitter());
        } break;
        case EEmitter: m_emitters.push_back(static_cast<Emitter *>(obj)); break;
        case ESampler:
            if (m_sampler) throw NoriException("There can only be one sampler per scene!");
            m_sampler = static_cast<Sampler *>(obj);
            break;
        case ECamera:
            if (m_camera) throw NoriException("There can only be one camera per scene!");
            m_camera = static_cast<Camera *>(obj);
            break;
        case EIntegrator:
            if (m_integrator) throw NoriException("There can only be one integrator per scene!");
            m_integrator = static_cast<Integrator *>(obj);
            break;
        case EMedium:
            if (m_ambientMedium) throw NoriException("There can only be one ambient medium per scene.");
            m_ambientMedium = static_cast<Medium *>(obj);
            break;
        case ERenderer: case EDenoiser: case EPixelSampler: m_other.push_back(obj); break;
        default:
            throw NoriException("Scene::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
    }
}

void Scene::activate() {
    if (!m_camera) throw NoriException("No camera was specified!");
    if (!m_sampler) m_sampler = static_cast<Sampler *>(NoriObjectFactory::createInstance("independent"));   /* scene.cpp:66-68 */
}

std::string Scene::toString() const {
    return NoriException::format("Scene[shapes=%d, emitters=%d]", (int) m_shapes.size(), (int) m_emitters.size());
}

bool Scene::flatten(FlatScene &out, std::string *why, bool buildTree) const {
    auto fail = [&](const std::string &msg) { if (why) *why = msg; return false; };
    out = FlatScene();
    if (!m_integrator || m_integrator->getB200IntegratorType() < 0)
        return fail("integrator '" + (m_integrator ? m_integrator->toString() : std::string("<none>")) + "' has no GPU kernel");
    for (Shape *s : m_shapes)
        if (auto *m = dynamic_cast<Mesh *>(s)) {
            out.any_normals |= m->hasNormals();
            out.any_uvs |= m->hasUVs();
            out.any_tangents |= m->hasNormals() && m->hasUVs();
        }
    for (Emitter *e : m_emitters) {
        nb_emitter rec;
        if (!e->getB200EmitterData(rec, out)) return fail("emitter '" + e->toString() + "' has no GPU kernel");
        out.emitters.push_back(rec);
    }
    for (size_t i = 0; i < m_shapes.size(); ++i) {
        const Shape *s = m_shapes[i];
        nb_shape rec; memset(&rec, 0, sizeof(rec));
        rec.bsdf = rec.emitter = rec.medium = -1;
        s->getB200Shape(rec, out);
        if (s->getBSDF()) {
            nb_bsdf b;
            if (!s->getBSDF()->getB200MaterialData(b, out)) return fail("bsdf '" + s->getBSDF()->toString() + "' has no GPU kernel");
            rec.bsdf = (int32_t) out.bsdfs.size();
            out.bsdfs.push_back(b);
        }
        if (s->getNormalMap()) {
            const int t = s->getNormalMap()->getB200Texture(out);
            if (t < 0 || !(out.textures[t].flags & NB_TEXTURE_NORMAL_MAP)) return fail("normal map '" + s->getNormalMap()->toString() + "' has no GPU kernel");
            rec.normal_map = (uint32_t) t + 1u;
        }
        if (s->getMedium()) {
            nb_medium m;
            if (!s->getMedium()->getB200MediumData(m, out)) return fail("medium '" + s->getMedium()->toString() + "' has no GPU kernel");
            rec.medium = (int32_t) out.media.size();
            out.media.push_back(m);
        }
        for (size_t e = 0; e < m_emitters.size(); ++e) {
            const bool surface = m_emitters[e] == s->getEmitter(), inside = m_emitters[e] == s->getMediumEmitter();
            if (!surface && !inside) continue;
            if (surface) rec.emitter = (int32_t) e;
            out.emitters[e].shape = (int32_t) i;
            if (out.emitters[e].type == NB_EMITTER_VOLUME) {             /* VolumeEmitter::cloneAndInit, volumelight.cpp:20-22 */
                if (!s->getMedium()) return fail("There is no shape or medium attached to this volume light.");
                out.emitters[e].position[0] = 1 / s->getVolume();        /* Shape::pdfVolume, shape.cpp:103-106 */
                const uint32_t off = (uint32_t) out.params.size();       /* Shape::sampleVolume's box, shape.cpp:97-101 */
                memcpy(&out.emitters[e].position[1], &off, 4);
                const BoundingBox3f &b = s->getBoundingBox();
                for (int k = 0; k < 3; ++k) out.params.push_back(b.min[k]);
                for (int k = 0; k < 3; ++k) out.params.push_back(b.max[k] - b.min[k]);
            }
        }
        out.shapes.push_back(rec);
    }
    /* pad optional vertex attributes to the full vertex count */
    const size_t nv = out.positions.size() / 3;
    if (out.any_normals) out.normals.resize(nv * 3, 0.f);
    if (out.any_uvs) out.uvs.resize(nv * 2, 0.f);
    if (out.any_tangents) { out.tangents.resize(nv * 3, 0.f); out.bitangents.resize(nv * 3, 0.f); }
    /* emitter selection CDF from lightProb (scene.cpp:180-184, DiscretePDF::normalize) */
    if (!m_emitters.empty()) {
        out.emitter_cdf.push_back(0.f);
        for (Emitter *e : m_emitters) out.emitter_cdf.push_back(out.emitter_cdf.back() + e->lightProb);
        const float sum = out.emitter_cdf.back();
        if (sum > 0) { const float n = 1.0f / sum; for (size_t i = 1; i < out.emitter_cdf.size(); ++i) out.emitter_cdf[i] *= n; out.emitter_cdf.back() = 1.0f; }
    }
    if (m_ambientMedium) {
        nb_medium m;
        if (!m_ambientMedium->getB200MediumData(m, out)) return fail("ambient medium '" + m_ambientMedium->toString() + "' has no GPU kernel");
        if (m.type != NB_MEDIUM_VACUUM) { out.ambient_medium = (int32_t) out.media.size(); out.media.push_back(m); }
    }
    m_camera->getB200Data(out.camera);
    memset(&out.render, 0, sizeof(out.render));
    out.render.integrator = (uint32_t) m_integrator->getB200IntegratorType();
    m_integrator->getB200IntegratorParams(out.render.integrator_params);
    const ReconstructionFilter *rf = m_camera->getReconstructionFilter();
    out.render.filter_radius = rf->getRadius();
    for (int i = 0; i < NB_FILTER_RESOLUTION; ++i)
        out.render.filter_table[i] = rf->eval((out.render.filter_radius * i) / NB_FILTER_RESOLUTION);
    out.render.filter_table[NB_FILTER_RESOLUTION] = 0.0f;
    out.spp = (uint32_t) m_sampler->getSampleCount();
    if (m_sampler->isAdaptive()) { out.adaptive[0] = 4; out.adaptive[1] = (uint32_t) m_sampler->getInitialUniform(); }   /* block.h:31 */
    out.bvh_nodes.clear(); out.bvh_indices.clear();
    if (buildTree) buildBVH(m_shapes, out.bvh_nodes, out.bvh_indices);
    return true;
}

nb_scene_desc FlatScene::desc() const {
    nb_scene_desc d;
    memset(&d, 0, sizeof(d));
    d.n_shapes = (uint32_t) shapes.size(); d.shapes = shapes.data();
    d.n_vertices = positions.size() / 3; d.positions = positions.data();
    d.normals = normals.empty() ? nullptr : normals.data();
    d.uvs = uvs.empty() ? nullptr : uvs.data();
    d.tangents = tangents.empty() ? nullptr : tangents.data();
    d.bitangents = bitangents.empty() ? nullptr : bitangents.data();
    d.n_faces = faces.size() / 3; d.faces = faces.data();
    d.n_bsdfs = (uint32_t) bsdfs.size(); d.bsdfs = bsdfs.data();
    d.n_emitters = (uint32_t) emitters.size(); d.emitters = emitters.data();
    d.emitter_cdf = emitter_cdf.empty() ? nullptr : emitter_cdf.data();
    d.n_area_cdf = area_cdf.size(); d.area_cdf = area_cdf.data();
    d.camera = camera;
    d.n_media = (uint32_t) media.size(); d.media = media.empty() ? nullptr : media.data();
    d.ambient_medium = ambient_medium;
    d.n_density = density.size(); d.density = density.empty() ? nullptr : density.data();
    d.n_textures = (uint32_t) textures.size(); d.textures = textures.empty() ? nullptr : textures.data();
    d.n_texels = texels.size() / 4; d.texels = texels.empty() ? nullptr : texels.data();
    d.n_params = params.size(); d.params = params.empty() ? nullptr : params.data();
    return d;
}

void FlatScene::writeNbs(const std::string &path) const {
    struct Sec { const char *name; uint32_t elem; const void *data; size_t count; };
    std::vector<Sec> secs = {
        { "shapes", sizeof(nb_shape), shapes.data(), shapes.size() }, { "positions", 4, positions.data(), positions.size() } };
    if (!normals.empty()) secs.push_back({ "normals", 4, normals.data(), normals.size() });
    if (!uvs.empty()) secs.push_back({ "uvs", 4, uvs.data(), uvs.size() });
    if (!tangents.empty()) { secs.push_back({ "tangents", 4, tangents.data(), tangents.size() }); secs.push_back({ "bitangents", 4, bitangents.data(), bitangents.size() }); }
    secs.push_back({ "faces", 4, faces.data(), faces.size() });
    secs.push_back({ "bsdfs", sizeof(nb_bsdf), bsdfs.data(), bsdfs.size() });
    secs.push_back({ "emitters", sizeof(nb_emitter), emitters.data(), emitters.size() });
    secs.push_back({ "emitter_cdf", 4, emitter_cdf.data(), emitter_cdf.size() });
    secs.push_back({ "area_cdf", 4, area_cdf.data(), area_cdf.size() });
    if (!media.empty() || ambient_medium >= 0) {
        secs.push_back({ "media", sizeof(nb_medium), media.data(), media.size() });
        secs.push_back({ "ambient", 4, &ambient_medium, 1 });
        if (!density.empty()) secs.push_back({ "density", 4, density.data(), density.size() });
    }
    if (!textures.empty()) secs.push_back({ "textures", sizeof(nb_texture), textures.data(), textures.size() });
    if (!texels.empty()) secs.push_back({ "texels", 4, texels.data(), texels.size() });
    if (!params.empty()) secs.push_back({ "params", 4, params.data(), params.size() });
    secs.push_back({ "camera", sizeof(nb_camera), &camera, 1 });
    secs.push_back({ "render", sizeof(nb_render_params), &render, 1 });
    secs.push_back({ "spp", 4, &spp, 1 });
    if (adaptive[0]) secs.push_back({ "adaptive", 4, adaptive, 2 });
    secs.push_back({ "bvh_nodes", sizeof(nb_bvh_node), bvh_nodes.data(), bvh_nodes.size() });
    secs.push_back({ "bvh_indices", 4, bvh_indices.data(), bvh_indices.size() });
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    const uint32_t hdr[2] = { NBS_MAGIC, (uint32_t) secs.size() };
    fwrite(hdr, 4, 2, f);
    uint64_t off = 8 + sizeof(nbs_section) * secs.size();
    std::vector<nbs_section> table(secs.size());
    for (size_t i = 0; i < secs.size(); ++i) {
        off = (off + 15) & ~(uint64_t) 15;
        memset(&table[i], 0, sizeof(nbs_section));
        strncpy(table[i].name, secs[i].name, 15);
        table[i].elem_size = secs[i].elem; table[i].count = secs[i].count; table[i].offset = off;
        off += (uint64_t) secs[i].elem * secs[i].count;
    }
    fwrite(table.data(), sizeof(nbs_section), table.size(), f);
    for (size_t i = 0; i < secs.size(); ++i) {
        for (uint64_t k = (uint64_t) ftell(f); k < table[i].offset; ++k) fputc(0, f);
        if (secs[i].count) fwrite(secs[i].data, secs[i].elem, secs[i].count, f);
    }
    fclose(f);
}

NORI_REGISTER_CLASS(Scene, "scene")

} // namespace nori
