/* oracle/ref_heterog.cpp — TEST INFRASTRUCTURE, never on the product path.
 *
 * The heterogeneous-medium row of the hot path (SURVEY.md 8a a30/a31, 8c "Volume path") has no
 * runnable reference: src/media/heterogmedium.cpp returns transmittance 0 by construction (:151),
 * mixes index- and world-space ray parameters (:120-129), samples through an accessor that is
 * destroyed before first use (src/textures/NvdbVolume.vdb.cpp:72-75), and the grid can only be
 * loaded through OpenVDB (absent here). PARITY UNPINNED: no test, golden or report number of the
 * reference fixes what a heterogeneous render should be.
 *
 * This file is the CPU restatement SURVEY.md 8c prescribes, linked into oracle/_ref/nori_ref in
 * place of src/media/heterogmedium.cpp and of NvdbVolume::loadFromFile:
 *   - it keeps the reference's structure (heterogmedium.cpp:105-174): clip the ray to the grid's
 *     worldBBox with nanovdb::Ray, delta tracking for the free path, ratio tracking for the
 *     transmittance, majorant max(0.001, densityScale*gridMax) * max(sigma_t), collision test
 *     density/gridMax against one uniform draw;
 *   - it uses the reference's vendored NanoVDB for everything NanoVDB does there: the grid
 *     (built with nanovdb::GridBuilder from a dense dump), nanovdb::Ray::intersects, and the
 *     27-tap SampleFromVoxels<.., 2> of NvdbVolume.h:39-41;
 *   - with the listed corrections: transmittance starts at 1; a ray that misses the box gives
 *     free path infinity / transmittance 1; the sampler owns its accessor (one per call, so the
 *     render threads do not share a cache); ray parameters are world distances throughout
 *     (points are mapped world->index one by one instead of marching an index-space ray); a free
 *     path that leaves the box without a real collision is infinity.
 * Grid input: a ".nbv" dense dump (magic "NBV1", i32 dims[3], i32 bbox_min[3], f64
 * index_to_world[12] row-major 3x4, f32 background, f32 values x-slowest) written by
 * libnorib200_host's own .vdb reader (nbh_vdb_to_nbv); a scene that names a .vdb is served from
 * $NORI_ORACLE_NBV_DIR/<stem>.nbv; a scene that names a .nvdb goes through the reference's own
 * nanovdb::io::readGrid, as NvdbVolume::loadNanoVdb does. */
#include <nori/medium.h>
#include <nori/sampler.h>
#include <nori/NvdbVolume.h>
#include <nanovdb/NanoVDB.h>
#include <nanovdb/util/Ray.h>
#include <nanovdb/util/GridBuilder.h>
#include <nanovdb/util/IO.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

NORI_NAMESPACE_BEGIN

void NvdbVolume::loadFromFile() {
    std::filesystem::path file = filename;
    if (file.extension() == ".vdb") {
        const char *dir = getenv("NORI_ORACLE_NBV_DIR");
        if (!dir) throw NoriException("oracle harness: %s needs OpenVDB; set NORI_ORACLE_NBV_DIR to a directory holding its .nbv dump", filename.string());
        file = std::filesystem::path(dir) / file.stem();
        file += ".nbv";
    }
    if (file.extension() == ".nvdb") {                       /* NvdbVolume::readGrid (NvdbVolume.vdb.cpp:42-61): grid 0 of the file */
        densityHandle = nanovdb::io::readGrid(file.string(), 0);
        densityGrid = densityHandle.grid<float>();
        if (!densityGrid) throw NoriException("GridHandle 0 does not contain a grid with value type float. (%s)", file.string());
        temperatureGrid = densityGrid;
        return;
    }
    FILE *f = fopen(file.string().c_str(), "rb");
    if (!f) throw NoriException("oracle harness: cannot open %s", file.string());
    char magic[4]; int32_t dims[3], org[3]; double M[12]; float background;
    bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "NBV1", 4) == 0 && fread(dims, 4, 3, f) == 3 && fread(org, 4, 3, f) == 3 &&
              fread(M, 8, 12, f) == 12 && fread(&background, 4, 1, f) == 1;
    std::vector<float> values;
    if (ok) { values.resize((size_t) dims[0] * dims[1] * dims[2]); ok = fread(values.data(), 4, values.size(), f) == values.size(); }
    fclose(f);
    if (!ok) throw NoriException("oracle harness: %s is not a complete .nbv file", file.string());
    nanovdb::GridBuilder<float> builder(background);
    {
        auto acc = builder.getAccessor();                    /* every voxel of the box is active, as in the file */
        size_t n = 0;
        for (int i = 0; i < dims[0]; ++i)
            for (int j = 0; j < dims[1]; ++j)
                for (int k = 0; k < dims[2]; ++k) acc.setValue(nanovdb::Coord(org[0] + i, org[1] + j, org[2] + k), values[n++]);
    }
    /* nanovdb::Map takes the 4x4 in OpenVDB's row-vector convention */
    const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], g = M[6], h = M[8], i = M[9], k = M[10];
    const double det = a * (e * k - g * i) - b * (d * k - g * h) + c * (d * i - e * h);
    const double inv[3][3] = { { (e * k - g * i) / det, (c * i - b * k) / det, (b * g - c * e) / det },
                               { (g * h - d * k) / det, (a * k - c * h) / det, (c * d - a * g) / det },
                               { (d * i - e * h) / det, (b * h - a * i) / det, (a * e - b * d) / det } };
    double mat[4][4] = {}, invMat[4][4] = {};
    for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 3; ++cc) { mat[cc][r] = M[r * 4 + cc]; invMat[cc][r] = inv[r][cc]; }
    for (int r = 0; r < 3; ++r) { mat[3][r] = M[r * 4 + 3]; invMat[3][r] = -M[r * 4 + 3]; }   /* Map keeps the translation, not its inverse image */
    mat[3][3] = invMat[3][3] = 1.0;
    nanovdb::Map map;
    map.set(mat, invMat, 1.0);
    densityHandle = builder.getHandle<nanovdb::HostBuffer>(map, "density", nanovdb::GridClass::FogVolume);
    densityGrid = densityHandle.grid<float>();
    if (!densityGrid) throw NoriException("oracle harness: grid build failed for %s", file.string());
    temperatureGrid = densityGrid;                           /* NvdbVolume.vdb.cpp:74-75 samples the density grid for both */
}

struct HeterogeneousMediumRestated : public Medium {
    explicit HeterogeneousMediumRestated(const PropertyList &propList) : Medium(propList) {
        m_densityScale = propList.getFloat("densityScale", 1.f);
        m_temperatureScale = propList.getFloat("temperatureScale", 0.f);
    }
    NoriObject *cloneAndInit() override {
        if (!m_volume) throw NoriException("HeterogeneousMedium requires a nvdb volume.");
        updateDerivedProperties();
        auto clone = new HeterogeneousMediumRestated(*this);
        clone->m_volume = static_cast<NvdbVolume *>(m_volume->cloneAndInit());
        Medium::cloneAndInit(clone);
        return clone;
    }
    void update(const NoriObject *guiObject) override {
        const auto *gui = static_cast<const HeterogeneousMediumRestated *>(guiObject);
        if (!gui->touched) return;
        gui->touched = false;
        m_densityScale = gui->m_densityScale;
        m_temperatureScale = gui->m_temperatureScale;
        m_volume->update(gui->m_volume);
        Medium::update(guiObject);
        m_densityMaxUnscaled = m_volume->densityGrid->tree().root().valueMax();
        m_densityMaxInvUnscaled = 1.f / m_densityMaxUnscaled;
        m_densityMax = std::max(0.001f, m_densityScale * m_densityMaxUnscaled);
        m_densityMaxInv = 1.f / m_densityMax;
    }
    ~HeterogeneousMediumRestated() override { delete m_volume; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EVolume)
            throw NoriException("HeterogeneousMedium::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        if (m_volume) throw NoriException("There is already an volume defined!");
        m_volume = static_cast<NvdbVolume *>(obj);
    }

    /* world-space [t0, t1] of the ray inside the grid's world box (heterogmedium.cpp:109-118) */
    bool clip(const Point3f &o, const Vector3f &d, float mint, float maxt, float &t0, float &t1) const {
        nanovdb::Ray<float> wRay(nanovdb::Vec3f(o.x(), o.y(), o.z()), nanovdb::Vec3f(d.x(), d.y(), d.z()), mint, maxt);
        return wRay.intersects(m_volume->densityGrid->worldBBox(), t0, t1);
    }
    float density(const Point3f &pWorld) const {
        const auto &grid = *m_volume->densityGrid;
        auto acc = grid.getAccessor();
        nanovdb::SampleFromVoxels<nanovdb::DefaultReadAccessor<float>, NvdbVolume::InterpolationOrder, false> sampler(acc);
        return sampler(grid.worldToIndexF(nanovdb::Vec3f(pWorld.x(), pWorld.y(), pWorld.z())));
    }
    float collisionProbability(float dens) const { return dens * m_densityScale * m_densityMaxInv; }   /* = dens / gridMax unless the 0.001 clamp is active */

    float sampleFreePath(MediumQueryRecord &mRec, Sampler &sampler) const override {
        float t, t1;
        if (!clip(mRec.ray.o, mRec.ray.d, mRec.ray.mint, mRec.ray.maxt, t, t1)) return INFINITY;
        while (true) {
            t += -std::log(sampler.next1D()) * m_densityMaxInv / m_sigma_t.maxCoeff();
            if (!(t < t1)) return INFINITY;
            if (collisionProbability(density(mRec.ray(t))) >= sampler.next1D()) return t;
        }
    }
    Color3f getTransmittance(const Vector3f &from, const Vector3f &to, const bool &, Sampler &sampler) const override {
        const float dist = (to - from).norm();
        if (!(dist > 0.f)) return Color3f(1.f);
        Ray3f ray(from, (to - from) / dist, Epsilon, dist);
        float t, t1;
        if (!clip(ray.o, ray.d, ray.mint, ray.maxt, t, t1)) return Color3f(1.f);
        float tr = 1.f;
        while (true) {
            t += -std::log(sampler.next1D()) * m_densityMaxInv / m_sigma_t.maxCoeff();
            if (!(t < t1)) break;
            tr *= 1.f - collisionProbability(density(ray(t)));
            if (tr <= 0.f) { tr = 0.f; break; }
        }
        return Color3f(tr);
    }
    std::string toString() const override { return "HeterogeneousMedium[restated]"; }
#ifdef NORI_USE_IMGUI
    NORI_OBJECT_IMGUI_NAME("Heterogeneous (restated)");
    bool getImGuiNodes() override { return touched; }
#endif
    float densityAt(const Point3f &p) const { return density(p); }
private:
    NvdbVolume *m_volume = nullptr;
    float m_densityScale, m_densityMaxUnscaled = 0.f, m_densityMaxInvUnscaled = 0.f, m_densityMax = 0.f, m_densityMaxInv = 0.f;
    float m_temperatureScale;
};

NORI_REGISTER_CLASS(HeterogeneousMediumRestated, "heterog");

/* for the harness' `medium` mode: NvdbVolume::getDensity at a world point */
float oracleHeterogDensity(const Medium *m, const Point3f &p, bool &isHeterog) {
    const auto *h = dynamic_cast<const HeterogeneousMediumRestated *>(m);
    isHeterog = h != nullptr;
    return h ? h->densityAt(p) : 0.f;
}

NORI_NAMESPACE_END
