/* plugins.h — the plugin base classes of the reference (BSDF, Emitter, Shape, Camera,
 * ReconstructionFilter, Integrator, Sampler, Scene) reduced to what the GPU backend needs from
 * them: construction from a PropertyList with the reference's defaults, the parent/child wiring,
 * and the flatten hooks that emit the C-ABI records. */
#pragma once
#include "nori.h"

namespace nori {

std::string resolvePath(const std::string &relative);      /* filesystem::resolver of the reference */
void setResolveRoot(const std::string &dir);

class Shape;
struct FlatScene;

/* include/nori/texture.h:35-78, Texture<Color3f> */
class Texture : public NoriObject {
public:
    EClassType getClassType() const override { return ETexture; }
    virtual Color3f eval(const Point2f &uv) const = 0;
    virtual unsigned getWidth() const { return 1; }
    virtual unsigned getHeight() const { return 1; }
    /* replaces getOptixTexture (texture.h:75): appends the record (and texels) to `flat`, returns its
     * index or -1 when the texture has no kernel */
    virtual int getB200Texture(FlatScene &flat) const = 0;
};

/* include/nori/bsdf.h:82-123 */
class BSDF : public NoriObject {
public:
    EClassType getClassType() const override { return EBSDF; }
    /* replaces getOptixMaterialData (bsdf.h:122); false = no kernel for this plugin */
    virtual bool getB200MaterialData(nb_bsdf &out, FlatScene &flat) const = 0;
};

/* include/nori/emitter.h:72-166 */
class Emitter : public NoriObject {
public:
    EClassType getClassType() const override { return EEmitter; }
    void setShape(Shape *s) { m_shape = s; }
    bool hasShape() const { return m_shape != nullptr; }
    const Shape *getShape() const { return m_shape; }
    virtual bool getB200EmitterData(nb_emitter &out, FlatScene &flat) const = 0;     /* replaces getOptixEmitterData (emitter.h:159) */
    virtual bool isEnvMap() const { return false; }
    float lightProb = 1.f;                                            /* emitter.h:151 */
protected:
    Shape *m_shape = nullptr;
    Color3f m_radiance;
};

/* include/nori/phase.h:28-41 */
class PhaseFunction : public NoriObject {
public:
    EClassType getClassType() const override { return EPhaseFunction; }
    virtual void getB200PhaseData(nb_medium &out) const = 0;          /* fills phase, phase_param */
};

/* include/nori/medium.h + src/media/medium.cpp */
class Medium : public NoriObject {
public:
    explicit Medium(const PropertyList &p);
    ~Medium() override;
    EClassType getClassType() const override { return EMedium; }
    void addChild(NoriObject *obj) override;            /* medium.cpp:50-72: one phase, one emitter */
    void activate() override;                           /* isotropic phase by default (medium.cpp:22-24) */
    /* replaces getOptixMediumData (medium.h:68); false = no GPU kernel for this medium */
    Emitter *getEmitter() const { return m_emitter; }
    virtual bool getB200MediumData(nb_medium &out, FlatScene &flat) const = 0;   /* heterogeneous media append their voxels to flat.density */
protected:
    Color3f m_sigma_a, m_sigma_s;                       /* normalized colour * intensity */
    PhaseFunction *m_phase = nullptr;
    Emitter *m_emitter = nullptr;                       /* volume emitter (volumelight.cpp) */
};

/* include/nori/shape.h:104-237 */
class Shape : public NoriObject {
public:
    ~Shape() override;
    EClassType getClassType() const override { return EShape; }
    void addChild(NoriObject *child) override;          /* src/shapes/shape.cpp:113-153 */
    void activate() override;                           /* default diffuse BSDF, shape.cpp:31-35 */
    virtual uint32_t getPrimitiveCount() const = 0;
    virtual BoundingBox3f getBoundingBox(uint32_t index) const = 0;
    virtual Point3f getCentroid(uint32_t index) const = 0;
    const BoundingBox3f &getBoundingBox() const { return m_bbox; }
    const BSDF *getBSDF() const { return m_bsdf; }
    const Medium *getMedium() const { return m_medium; }
    Emitter *getEmitter() const { return m_emitter; }
    const Texture *getNormalMap() const { return m_normalMap; }
    Emitter *getMediumEmitter() const { return m_medium ? m_medium->getEmitter() : nullptr; }       /* shape.h */
    virtual float getVolume() const {                   /* Shape::updateVolume, shape.cpp:108-111: the bounding box's */
        return (m_bbox.max[0] - m_bbox.min[0]) * (m_bbox.max[1] - m_bbox.min[1]) * (m_bbox.max[2] - m_bbox.min[2]);
    }
    bool isEmitter() const { return m_emitter != nullptr; }
    /* replaces getOptixBuildInput / getOptixHitgroupRecords (shape.h:194-201): appends this
     * shape's geometry to the flat arrays and fills the geometric part of the record */
    virtual void getB200Shape(nb_shape &rec, struct FlatScene &flat) const = 0;
protected:
    BSDF *m_bsdf = nullptr;
    Emitter *m_emitter = nullptr;
    Medium *m_medium = nullptr;
    Texture *m_normalMap = nullptr;                     /* shape.h:226: Texture<Normal3f>, a png_texture named "normal" */
    BoundingBox3f m_bbox;
};

/* include/nori/mesh.h + src/shapes/mesh.cpp */
class Mesh : public Shape {
public:
    uint32_t getPrimitiveCount() const override { return (uint32_t) (m_F.size() / 3); }
    BoundingBox3f getBoundingBox(uint32_t index) const override;
    Point3f getCentroid(uint32_t index) const override;
    float surfaceArea(uint32_t index) const;
    void activate() override;                           /* area DiscretePDF, mesh.cpp:35-48 */
    void getB200Shape(nb_shape &rec, FlatScene &flat) const override;
    std::string toString() const override;
    bool hasNormals() const { return !m_N.empty(); }
    bool hasUVs() const { return !m_UV.empty(); }
    Point3f vertex(uint32_t i) const { return Point3f(m_V[3 * i], m_V[3 * i + 1], m_V[3 * i + 2]); }
protected:
    std::string m_name;
    std::vector<float> m_V, m_N, m_UV, m_T, m_BT;       /* 3,3,2,3,3 floats per vertex */
    std::vector<uint32_t> m_F;
    std::vector<float> m_cdf;                           /* DiscretePDF::m_cdf */
    float m_cdfNormalization = 0.f;
};

/* include/nori/camera.h, src/cameras/perspective.cpp, rfilter.cpp */
class ReconstructionFilter : public NoriObject {
public:
    EClassType getClassType() const override { return EReconstructionFilter; }
    float getRadius() const { return m_radius; }
    virtual float eval(float x) const = 0;
protected:
    float m_radius = 0.f;
};

class Camera : public NoriObject {
public:
    ~Camera() override { delete m_rfilter; }
    EClassType getClassType() const override { return ECamera; }
    const Vector2i &getOutputSize() const { return m_outputSize; }
    const ReconstructionFilter *getReconstructionFilter() const { return m_rfilter; }
    virtual void getB200Data(nb_camera &out) const = 0;               /* replaces getOptixData (camera.h:86) */
protected:
    Vector2i m_outputSize;
    ReconstructionFilter *m_rfilter = nullptr;
};

class Integrator : public NoriObject {
public:
    EClassType getClassType() const override { return EIntegrator; }
    virtual int getB200IntegratorType() const = 0;                    /* replaces getOptixIntegratorType (integrator.h:91); -1 = none */
    virtual void getB200IntegratorParams(float out[3]) const { (void) out; }   /* nb_render_params.integrator_params */
};

class Sampler : public NoriObject {
public:
    EClassType getClassType() const override { return ESampler; }
    size_t getSampleCount() const { return m_sampleCount; }
    virtual bool isAdaptive() const { return false; }           /* sampler.h:117 */
    virtual int getInitialUniform() const { return 0; }
protected:
    size_t m_sampleCount = 1;
};

/* Flat arrays behind nb_scene_desc, plus the BVH and render settings (= one .nbs file) */
struct FlatScene {
    std::vector<nb_shape> shapes;
    std::vector<float> positions, normals, uvs, tangents, bitangents, area_cdf, emitter_cdf;
    std::vector<uint32_t> faces;
    std::vector<nb_bsdf> bsdfs;
    std::vector<nb_emitter> emitters;
    std::vector<nb_medium> media;
    std::vector<float> density;                         /* voxel arrays of heterogeneous media */
    std::vector<nb_texture> textures;
    std::vector<float> texels;                          /* RGBA */
    std::vector<float> params;                          /* nb_scene_desc.params */
    int32_t ambient_medium = -1;
    nb_camera camera;
    nb_render_params render;
    uint32_t spp = 1;
    uint32_t adaptive[2] = { 0, 0 };          /* `adaptive` sampler: { block size (NORI_BLOCK_SIZE_ADAPTIVE), initialUniform }; { 0, 0 } otherwise */
    std::vector<nb_bvh_node> bvh_nodes;
    std::vector<uint32_t> bvh_indices;
    bool any_normals = false, any_uvs = false, any_tangents = false;
    nb_scene_desc desc() const;
    void writeNbs(const std::string &path) const;
};

/* include/nori/scene.h + src/utils/scene.cpp */
class Scene : public NoriObject {
public:
    explicit Scene(const PropertyList &);
    ~Scene() override;
    EClassType getClassType() const override { return EScene; }
    void addChild(NoriObject *obj) override;            /* scene.cpp:204-266 */
    void activate() override;                           /* defaults (scene.cpp:59-151) */
    std::string toString() const override;
    const std::vector<Shape *> &getShapes() const { return m_shapes; }
    const std::vector<Emitter *> &getLights() const { return m_emitters; }
    const Camera *getCamera() const { return m_camera; }
    const Integrator *getIntegrator() const { return m_integrator; }
    const Sampler *getSampler() const { return m_sampler; }
    /* OptixState::preRender's gathering loop (OptixState.render.cpp:19-72): every plugin's hook,
     * then the BVH (built here with the reference's SAH algorithm, src/utils/bvh.cpp:100-380).
     * Returns false when some plugin has no GPU kernel (caller keeps its CPU renderer). */
    /* buildTree = false: no host SAH build (bvh_nodes / bvh_indices stay empty); the caller builds the tree on the device */
    bool flatten(FlatScene &out, std::string *why = nullptr, bool buildTree = true) const;
private:
    std::vector<Shape *> m_shapes;
    std::vector<Emitter *> m_emitters;
    Camera *m_camera = nullptr;
    Integrator *m_integrator = nullptr;
    Sampler *m_sampler = nullptr;
    Medium *m_ambientMedium = nullptr;                  /* scene.cpp:248-252; vacuum when absent */
    std::vector<NoriObject *> m_other;
};

/* src/utils/bvh.cpp:100-380 restated: SAH build over all primitives of the shapes (binned for
 * >= 32 primitives, exact sweep below), node layout and numbering as BVH::BVHNode, compacted. */
void buildBVH(const std::vector<Shape *> &shapes, std::vector<nb_bvh_node> &nodes, std::vector<uint32_t> &indices);

} // namespace nori
