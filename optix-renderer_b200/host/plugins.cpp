/* plugins.cpp — BSDF / emitter / camera / filter / integrator / sampler / sphere plugins.
 * Each constructor reads the properties the reference's constructor reads, with its defaults
 * (file:line beside each class); each flatten hook fills the matching C-ABI record. */
#include "plugins.h"
#include "vdb.h"
#include <cstring>

namespace nori {

static thread_local std::string g_resolveRoot;
void setResolveRoot(const std::string &dir) { g_resolveRoot = dir; }
std::string resolvePath(const std::string &rel) {
    if (rel.empty() || rel[0] == '/' || g_resolveRoot.empty()) return rel;
    return g_resolveRoot + "/" + rel;
}

/* ---- BSDFs ------------------------------------------------------------------------------------- */
class Diffuse : public BSDF {                       /* src/bsdf/diffuse.cpp:31-91: constant albedo (default 0.5) or an "albedo" texture */
public:
    explicit Diffuse(const PropertyList &p) : m_albedo(p.getColor("albedo", Color3f(0.5f))), m_hasConstant(p.has("albedo")) {}
    ~Diffuse() override { delete m_texture; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != ETexture) throw NoriException("Diffuse::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        if (obj->getIdName() != "albedo") throw NoriException("The name of this texture does not match any field!");
        if (m_texture || m_hasConstant) throw NoriException("There is already an albedo defined!");
        m_texture = static_cast<Texture *>(obj);
    }
    bool getB200MaterialData(nb_bsdf &o, FlatScene &flat) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_BSDF_DIFFUSE;
        for (int i = 0; i < 3; ++i) o.albedo[i] = m_albedo[i];
        if (m_texture) {
            const int t = m_texture->getB200Texture(flat);
            if (t < 0) return false;
            o.type |= (uint32_t) (t + 1) << 8;
        }
        return true;
    }
    std::string toString() const override { return NoriException::format("Diffuse[albedo=[%f, %f, %f]]", m_albedo[0], m_albedo[1], m_albedo[2]); }
private:
    Color3f m_albedo;
    bool m_hasConstant;
    Texture *m_texture = nullptr;
};
NORI_REGISTER_CLASS(Diffuse, "diffuse")

class Mirror : public BSDF {                        /* src/bsdf/mirror.cpp */
public:
    explicit Mirror(const PropertyList &) {}
    bool getB200MaterialData(nb_bsdf &o, FlatScene &) const override { memset(&o, 0, sizeof(o)); o.type = NB_BSDF_MIRROR; return true; }
    std::string toString() const override { return "Mirror[]"; }
};
NORI_REGISTER_CLASS(Mirror, "mirror")

class Dielectric : public BSDF {                    /* src/bsdf/dielectric.cpp:28-35 */
public:
    explicit Dielectric(const PropertyList &p)
        : m_intIOR(p.getFloat("intIOR", 1.5046f)), m_extIOR(p.getFloat("extIOR", 1.000277f)) {}
    bool getB200MaterialData(nb_bsdf &o, FlatScene &) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_BSDF_DIELECTRIC; o.int_ior = m_intIOR; o.ext_ior = m_extIOR;
        return true;
    }
    std::string toString() const override { return NoriException::format("Dielectric[intIOR=%f, extIOR=%f]", m_intIOR, m_extIOR); }
private:
    float m_intIOR, m_extIOR;
};
NORI_REGISTER_CLASS(Dielectric, "dielectric")

class Microfacet : public BSDF {                    /* src/bsdf/microfacet.cpp:33-53 */
public:
    explicit Microfacet(const PropertyList &p)
        : m_alpha(p.getFloat("alpha", 0.1f)), m_intIOR(p.getFloat("intIOR", 1.5046f)),
          m_extIOR(p.getFloat("extIOR", 1.000277f)), m_kd(p.getColor("kd", Color3f(0.5f))) {
        m_ks = 1 - m_kd.maxCoeff();
    }
    bool getB200MaterialData(nb_bsdf &o, FlatScene &) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_BSDF_MICROFACET; o.alpha = m_alpha; o.int_ior = m_intIOR; o.ext_ior = m_extIOR; o.ks = m_ks;
        for (int i = 0; i < 3; ++i) o.albedo[i] = m_kd[i];
        return true;
    }
    std::string toString() const override { return NoriException::format("Microfacet[alpha=%f, intIOR=%f, extIOR=%f, ks=%f]", m_alpha, m_intIOR, m_extIOR, m_ks); }
private:
    float m_alpha, m_intIOR, m_extIOR, m_ks;
    Color3f m_kd;
};
NORI_REGISTER_CLASS(Microfacet, "microfacet")

/* src/bsdf/disney.cpp:19-109: ten scalars clamped to [0, 1], base colour constant (default 0.5) or an "albedo" texture */
class Disney : public BSDF {
public:
    explicit Disney(const PropertyList &p) : m_albedo(p.getColor("albedo", Color3f(0.5f))), m_hasConstant(p.has("albedo")) {
        auto c = [&](const char *n, float d) { return std::min(std::max(p.getFloat(n, d), 0.f), 1.f); };
        m_p[0] = c("metallic", 0.f); m_p[1] = c("subsurface", 0.f); m_p[2] = c("specular", 0.5f); m_p[3] = c("roughness", 0.5f);
        m_p[4] = c("specularTint", 0.f); m_p[5] = c("anisotropic", 0.f); m_p[6] = c("sheen", 0.f); m_p[7] = c("sheenTint", 0.5f);
        m_p[8] = c("clearcoat", 0.f); m_p[9] = c("clearcoatGloss", 1.f);
    }
    ~Disney() override { delete m_texture; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != ETexture) throw NoriException("Disney::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        if (obj->getIdName() != "albedo") throw NoriException("The name of this texture does not match any field!");
        if (m_texture || m_hasConstant) throw NoriException("There is already an albedo defined!");
        m_texture = static_cast<Texture *>(obj);
    }
    bool getB200MaterialData(nb_bsdf &o, FlatScene &flat) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_BSDF_DISNEY;
        for (int i = 0; i < 3; ++i) o.albedo[i] = m_albedo[i];
        if (m_texture) {
            const int t = m_texture->getB200Texture(flat);
            if (t < 0) return false;
            o.type |= (uint32_t) (t + 1) << 8;
        }
        const uint32_t off = (uint32_t) flat.params.size();
        memcpy(&o.alpha, &off, 4);
        flat.params.insert(flat.params.end(), m_p, m_p + 10);
        return true;
    }
    std::string toString() const override { return NoriException::format("Disney[metallic=%f, roughness=%f]", m_p[0], m_p[3]); }
private:
    Color3f m_albedo;
    bool m_hasConstant;
    Texture *m_texture = nullptr;
    float m_p[10];
};
NORI_REGISTER_CLASS(Disney, "disney")

/* ---- emitters ---------------------------------------------------------------------------------- */
class AreaEmitter : public Emitter {                /* src/emitters/arealight.cpp:27-33 */
public:
    explicit AreaEmitter(const PropertyList &p) { m_radiance = p.getColor("radiance"); lightProb = p.getFloat("lightWeight", 1.f); }
    bool getB200EmitterData(nb_emitter &o, FlatScene &) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_AREA; o.shape = -1;
        for (int i = 0; i < 3; ++i) o.radiance[i] = m_radiance[i];
        return true;
    }
    std::string toString() const override { return NoriException::format("AreaLight[radiance=[%f, %f, %f]]", m_radiance[0], m_radiance[1], m_radiance[2]); }
};
NORI_REGISTER_CLASS(AreaEmitter, "area")

class PointLight : public Emitter {                 /* src/emitters/pointlight.cpp:14-45 */
public:
    explicit PointLight(const PropertyList &p) {
        m_power = p.getColor("power"); m_position = p.getPoint3("position"); lightProb = p.getFloat("lightWeight", 1.f);
        m_radiance = m_power / (float) (4 * M_PI);                       /* pointlight.cpp:44 */
    }
    bool getB200EmitterData(nb_emitter &o, FlatScene &) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_POINT; o.shape = -1;
        for (int i = 0; i < 3; ++i) { o.radiance[i] = m_radiance[i]; o.position[i] = m_position[i]; }
        return true;
    }
    std::string toString() const override { return "PointLight[]"; }
private:
    Color3f m_power;
    Point3f m_position;
};
NORI_REGISTER_CLASS(PointLight, "point")

/* src/emitters/volumelight.cpp:14-77: constant radiance inside the volume of the shape that owns it (directly, or
 * through its medium); sampled uniformly over Shape::sampleVolume (sphere: the ball; mesh: its bounding box) */
class VolumeEmitter : public Emitter {
public:
    explicit VolumeEmitter(const PropertyList &p) { m_radiance = p.getColor("radiance"); lightProb = p.getFloat("lightWeight", 1.f); }
    bool getB200EmitterData(nb_emitter &o, FlatScene &) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_VOLUME; o.shape = -1;                        /* owner and volume are filled in by Scene::flatten */
        for (int i = 0; i < 3; ++i) o.radiance[i] = m_radiance[i];
        return true;
    }
    std::string toString() const override { return "VolumeEmitter[]"; }
};
NORI_REGISTER_CLASS(VolumeEmitter, "volumelight")

class SpotLight : public Emitter {                  /* src/emitters/spotlight.cpp:14-24, 48-49 */
public:
    explicit SpotLight(const PropertyList &p) {
        m_position = p.getPoint3("position", Point3f());
        m_direction = p.getVector3("direction", Vector3f()).normalized();
        m_power = p.getColor("power", Color3f());
        m_falloffStart = p.getFloat("falloffstart");
        m_totalWidth = p.getFloat("totalwidth");
        lightProb = p.getFloat("lightWeight", 1.f);
        m_radiance = m_power / 2.f / (float) M_PI;                       /* spotlight.cpp:49, set by the first update() */
    }
    bool getB200EmitterData(nb_emitter &o, FlatScene &flat) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_SPOT; o.shape = (int32_t) flat.params.size();
        for (int i = 0; i < 3; ++i) { o.radiance[i] = m_radiance[i]; o.position[i] = m_position[i]; flat.params.push_back(m_direction[i]); }
        flat.params.push_back(std::cos(refDegToRad(m_falloffStart)));    /* spotlight.cpp:65-66, 169-170 */
        flat.params.push_back(std::cos(refDegToRad(m_totalWidth / 2.f)));
        return true;
    }
    std::string toString() const override { return "Spotlight[]"; }
    static float refDegToRad(float v) { return (float) (v * (M_PI / 180.0f)); }   /* common.h:218: the product is formed in double */
private:
    Point3f m_position;
    Vector3f m_direction;
    Color3f m_power;
    float m_falloffStart, m_totalWidth;
};
NORI_REGISTER_CLASS(SpotLight, "spot")

class DirectionalLight : public Emitter {           /* src/emitters/directionalLight.cpp:31-38 */
public:
    explicit DirectionalLight(const PropertyList &p) {
        m_direction = p.getVector3("direction", Vector3f(0, 0, 1)).normalized();
        m_radiance = p.getColor("radiance", Color3f(0.f));
        m_angle = p.getFloat("angle", 1.f);
    }
    bool getB200EmitterData(nb_emitter &o, FlatScene &flat) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_DIRECTIONAL; o.shape = (int32_t) flat.params.size();
        for (int i = 0; i < 3; ++i) { o.radiance[i] = m_radiance[i]; flat.params.push_back(m_direction[i]); }
        flat.params.push_back((float) std::cos((double) SpotLight::refDegToRad(m_angle)));   /* cos(degToRad(m_angle)): the unqualified cos(float) is ::cos(double) */
        return true;
    }
    std::string toString() const override { return "DirectionalLight[]"; }
private:
    Vector3f m_direction;
    float m_angle;
};
NORI_REGISTER_CLASS(DirectionalLight, "directional")

/* ---- reconstruction filters (src/cameras/rfilter.cpp) -------------------------------------------- */
class GaussianFilter : public ReconstructionFilter {                     /* rfilter.cpp:30-47 */
public:
    explicit GaussianFilter(const PropertyList &p) { m_radius = p.getFloat("radius", 2.0f); m_stddev = p.getFloat("stddev", 0.5f); }
    float eval(float x) const override {
        float alpha = -1.0f / (2.0f * m_stddev * m_stddev);
        return std::max(0.0f, std::exp(alpha * x * x) - std::exp(alpha * m_radius * m_radius));
    }
    std::string toString() const override { return NoriException::format("GaussianFilter[radius=%f, stddev=%f]", m_radius, m_stddev); }
private:
    float m_stddev;
};
NORI_REGISTER_CLASS(GaussianFilter, "gaussian")

class MitchellNetravaliFilter : public ReconstructionFilter {            /* rfilter.cpp:80-113 */
public:
    explicit MitchellNetravaliFilter(const PropertyList &p) {
        m_radius = p.getFloat("radius", 2.0f); m_B = p.getFloat("B", 1.0f / 3.0f); m_C = p.getFloat("C", 1.0f / 3.0f);
    }
    float eval(float x) const override {
        x = std::abs(2.0f * x / m_radius);
        float x2 = x * x, x3 = x2 * x;
        if (x < 1) return 1.0f / 6.0f * ((12 - 9 * m_B - 6 * m_C) * x3 + (-18 + 12 * m_B + 6 * m_C) * x2 + (6 - 2 * m_B));
        if (x < 2) return 1.0f / 6.0f * ((-m_B - 6 * m_C) * x3 + (6 * m_B + 30 * m_C) * x2 + (-12 * m_B - 48 * m_C) * x + (8 * m_B + 24 * m_C));
        return 0.0f;
    }
    std::string toString() const override { return NoriException::format("MitchellNetravaliFilter[radius=%f, B=%f, C=%f]", m_radius, m_B, m_C); }
private:
    float m_B, m_C;
};
NORI_REGISTER_CLASS(MitchellNetravaliFilter, "mitchell")

class TentFilter : public ReconstructionFilter {                         /* rfilter.cpp:129-141 */
public:
    explicit TentFilter(const PropertyList &) { m_radius = 1.0f; }
    float eval(float x) const override { return std::max(0.0f, 1.0f - std::abs(x)); }
    std::string toString() const override { return "TentFilter[]"; }
};
NORI_REGISTER_CLASS(TentFilter, "tent")

class BoxFilter : public ReconstructionFilter {                          /* rfilter.cpp:150-163 */
public:
    explicit BoxFilter(const PropertyList &) { m_radius = 0.5f; }
    float eval(float) const override { return 1.0f; }
    std::string toString() const override { return "BoxFilter[]"; }
};
NORI_REGISTER_CLASS(BoxFilter, "box")

/* ---- camera (src/cameras/perspective.cpp:10-96) --------------------------------------------------- */
class PerspectiveCamera : public Camera {
public:
    explicit PerspectiveCamera(const PropertyList &p) {
        m_outputSize = Vector2i(p.getInteger("width", 1280), p.getInteger("height", 720));
        m_cameraToWorld = p.getTransform("toWorld", Transform());
        m_fov = p.getFloat("fov", 30.0f);
        m_nearClip = p.getFloat("nearClip", 1e-4f);
        m_farClip = p.getFloat("farClip", 1e4f);
        m_focalDistance = p.getFloat("focalDistance", 10.f);
        m_fstop = p.getFloat("fstop", 0.f);
        m_lensRadius = p.getFloat("lensRadius", 0.f);
    }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EReconstructionFilter) return Camera::addChild(obj);
        if (m_rfilter) throw NoriException("Camera: tried to register multiple reconstruction filters!");
        m_rfilter = static_cast<ReconstructionFilter *>(obj);
    }
    void activate() override {                                           /* perspective.cpp:33-46, 67-95 */
        if (!m_rfilter) m_rfilter = static_cast<ReconstructionFilter *>(NoriObjectFactory::createInstance("gaussian"));
        if (m_fstop == 0.f) m_fstop = m_focalDistance / m_lensRadius; else m_lensRadius = m_focalDistance / m_fstop;
        const float aspect = m_outputSize.x() / (float) m_outputSize.y();
        const float recip = 1.0f / (m_farClip - m_nearClip), cot = 1.0f / std::tan(degToRad(m_fov / 2.0f));
        Matrix4f persp;
        memset(persp.m, 0, sizeof(persp.m));
        persp(0, 0) = cot; persp(1, 1) = cot; persp(2, 2) = m_farClip * recip; persp(2, 3) = -m_nearClip * m_farClip * recip; persp(3, 2) = 1;
        Matrix4f scale = Matrix4f::identity(), trans = Matrix4f::identity();
        scale(0, 0) = 0.5f; scale(1, 1) = -0.5f * aspect;
        trans(0, 3) = 1.0f; trans(1, 3) = -1.0f / aspect;
        m_sampleToCamera = Transform((scale * trans * persp).inverse());
    }
    void getB200Data(nb_camera &o) const override {
        memset(&o, 0, sizeof(o));
        memcpy(o.sample_to_camera, m_sampleToCamera.m.m, sizeof(o.sample_to_camera));
        memcpy(o.camera_to_world, m_cameraToWorld.m.m, sizeof(o.camera_to_world));
        o.inv_output_size[0] = 1.0f / (float) m_outputSize.x(); o.inv_output_size[1] = 1.0f / (float) m_outputSize.y();
        o.near_clip = m_nearClip; o.far_clip = m_farClip; o.lens_radius = m_lensRadius; o.focal_distance = m_focalDistance;
        o.width = (uint32_t) m_outputSize.x(); o.height = (uint32_t) m_outputSize.y();
    }
    std::string toString() const override { return NoriException::format("PerspectiveCamera[fov=%f, %dx%d]", m_fov, m_outputSize.x(), m_outputSize.y()); }
private:
    float m_fov, m_nearClip, m_farClip, m_focalDistance, m_fstop, m_lensRadius;
    Transform m_cameraToWorld, m_sampleToCamera;
};
NORI_REGISTER_CLASS(PerspectiveCamera, "perspective")

/* ---- integrators: only the tag crosses the boundary (integrator.h:91) ------------------------------ */
#define NB_INTEGRATOR_CLASS(cls, name, tag) \
    class cls : public Integrator { public: explicit cls(const PropertyList &) {} \
        int getB200IntegratorType() const override { return tag; } \
        std::string toString() const override { return #cls "[]"; } }; \
    NORI_REGISTER_CLASS(cls, name)
NB_INTEGRATOR_CLASS(PathMISIntegrator, "path_mis", NB_INTEGRATOR_PATH_MIS)
NB_INTEGRATOR_CLASS(PathMATSIntegrator, "path_mats", NB_INTEGRATOR_PATH_MATS)
NB_INTEGRATOR_CLASS(DirectEMSIntegrator, "direct_ems", NB_INTEGRATOR_DIRECT_EMS)
NB_INTEGRATOR_CLASS(DirectMATSIntegrator, "direct_mats", NB_INTEGRATOR_DIRECT_MATS)
NB_INTEGRATOR_CLASS(DirectMISIntegrator, "direct_mis", NB_INTEGRATOR_DIRECT_MIS)
NB_INTEGRATOR_CLASS(DirectIntegrator, "direct", NB_INTEGRATOR_DIRECT)
class NormalIntegrator : public Integrator {        /* src/integrators/normals.cpp:11-13 */
public:
    explicit NormalIntegrator(const PropertyList &p) : m_direction(p.getPoint3("direction", Point3f(0, 0, 1))) {}
    int getB200IntegratorType() const override { return NB_INTEGRATOR_NORMALS; }
    void getB200IntegratorParams(float out[3]) const override { for (int i = 0; i < 3; ++i) out[i] = m_direction[i]; }
    std::string toString() const override { return "NormalIntegrator[]"; }
private:
    Vector3f m_direction;
};
NORI_REGISTER_CLASS(NormalIntegrator, "normals")
class AverageVisibility : public Integrator {       /* src/integrators/av.cpp:12-15 */
public:
    explicit AverageVisibility(const PropertyList &p) : m_length(p.getFloat("length")) {}
    int getB200IntegratorType() const override { return NB_INTEGRATOR_AV; }
    void getB200IntegratorParams(float out[3]) const override { out[0] = m_length; }
    std::string toString() const override { return NoriException::format("AverageVisibilityIntegrator[length=%f]", m_length); }
private:
    float m_length;
};
NORI_REGISTER_CLASS(AverageVisibility, "av")
NB_INTEGRATOR_CLASS(PathVolMATSIntegrator, "path_vol_mats", NB_INTEGRATOR_PATH_VOL_MATS)
NB_INTEGRATOR_CLASS(PathVolMISIntegrator, "path_vol_mis", NB_INTEGRATOR_PATH_VOL_MIS)
NB_INTEGRATOR_CLASS(PhotonMapper, "photonmapper", -1)
NB_INTEGRATOR_CLASS(PreviewIntegrator, "preview", NB_INTEGRATOR_PREVIEW)

class Independent : public Sampler {                /* src/samplers/independent.cpp:35-39 */
public:
    explicit Independent(const PropertyList &p) { m_sampleCount = (size_t) p.getInteger("sampleCount", 1); }
    std::string toString() const override { return NoriException::format("Independent[sampleCount=%d]", (int) m_sampleCount); }
};
NORI_REGISTER_CLASS(Independent, "independent")

/* src/samplers/adaptive.cpp:18-23: the variance feedback loop of render.cpp:323-336 runs on the device
 * (nb_render_adaptive_round, one call per round of the sample loop: B200State::renderAdaptiveRound) */
class AdaptiveSampler : public Sampler {
public:
    explicit AdaptiveSampler(const PropertyList &p) {
        m_sampleCount = (size_t) p.getInteger("sampleCount", 1);
        m_initialUniform = std::min(std::max(p.getInteger("initialUniform", 2), 1), (int) m_sampleCount);
    }
    bool isAdaptive() const override { return true; }
    int getInitialUniform() const override { return m_initialUniform; }
    std::string toString() const override { return NoriException::format("AdaptiveSampler[\n  sampleCount = %d,\n  initialUniform = %d\n  ]", (int) m_sampleCount, m_initialUniform); }
private:
    int m_initialUniform;
};
NORI_REGISTER_CLASS(AdaptiveSampler, "adaptive")

/* src/denoiser/simple.cpp: a CPU post-process of the finished image, outside the GPU path; accepted so its scenes load */
class SimpleDenoiser : public NoriObject {
public:
    explicit SimpleDenoiser(const PropertyList &p) { (void) p.getFloat("sigma_d", 0.f); (void) p.getFloat("sigma_vr", 0.6f); (void) p.getInteger("range", 1); (void) p.getInteger("amount", 1); }
    EClassType getClassType() const override { return EDenoiser; }
    std::string toString() const override { return "SimpleDenoiser[]"; }
};
NORI_REGISTER_CLASS(SimpleDenoiser, "simple")

/* the GPU side-car's own settings object (OptixRenderer, include/nori/optix/OptixRenderer.cpp:9-14) */
class B200Renderer : public NoriObject {
public:
    explicit B200Renderer(const PropertyList &p) : samplesPerLaunch(p.getInteger("samplesPerLaunch", 16)) {}
    EClassType getClassType() const override { return ERenderer; }
    std::string toString() const override { return NoriException::format("B200Renderer[samplesPerLaunch=%d]", samplesPerLaunch); }
    int samplesPerLaunch;
};
NORI_REGISTER_CLASS(B200Renderer, "b200")
class OptixAlias : public B200Renderer { public: explicit OptixAlias(const PropertyList &p) : B200Renderer(p) {} };
NORI_REGISTER_CLASS(OptixAlias, "optix")

/* ---- Shape base (src/shapes/shape.cpp) --------------------------------------------------------------- */
/* ---- phase functions (src/bsdf/isophase.cpp, anisophase.cpp, schlickphase.cpp) ---------------------------- */
class IsoPhase : public PhaseFunction {
public:
    explicit IsoPhase(const PropertyList &) {}
    void getB200PhaseData(nb_medium &out) const override { out.phase = NB_PHASE_ISOTROPIC; out.phase_param = 0.f; }
    std::string toString() const override { return "IsoPhase[]"; }
};
NORI_REGISTER_CLASS(IsoPhase, "isophase")

class AnisoPhase : public PhaseFunction {           /* anisophase.cpp:26-29: g default 0 */
public:
    explicit AnisoPhase(const PropertyList &p) : m_g(p.getFloat("g", 0.f)) {}
    void getB200PhaseData(nb_medium &out) const override { out.phase = NB_PHASE_HENYEY_GREENSTEIN; out.phase_param = m_g; }
    std::string toString() const override { return NoriException::format("AnisoPhase[g=%f]", m_g); }
private:
    float m_g;
};
NORI_REGISTER_CLASS(AnisoPhase, "anisophase")

class SchlickPhase : public PhaseFunction {         /* schlickphase.cpp:26-30: k = 1.55 g - 0.55 g^3 */
public:
    explicit SchlickPhase(const PropertyList &p) : m_g(p.getFloat("g", 0.f)) { m_k = 1.55f * m_g - 0.55f * std::pow(m_g, 3); }
    void getB200PhaseData(nb_medium &out) const override { out.phase = NB_PHASE_SCHLICK; out.phase_param = m_k; }
    std::string toString() const override { return NoriException::format("SchlickPhase[g=%f]", m_g); }
private:
    float m_g, m_k;
};
NORI_REGISTER_CLASS(SchlickPhase, "schlick")          /* the reference's name (schlickphase.cpp:81) */

/* ---- media (src/media/medium.cpp, homogmedium.cpp, vacuum.cpp, heterogmedium.cpp) --------------------------- */
Medium::Medium(const PropertyList &p) {             /* medium.cpp:10-17 + updateDerivedProperties :112-117 */
    m_sigma_a = p.getColor("sigma_a", Color3f(0.5f)) * p.getFloat("sigma_a_intensity", 1.f);
    m_sigma_s = p.getColor("sigma_s", Color3f(0.f)) * p.getFloat("sigma_s_intensity", 1.f);
}
Medium::~Medium() { delete m_phase; delete m_emitter; }
void Medium::addChild(NoriObject *obj) {
    switch (obj->getClassType()) {
        case EPhaseFunction:
            if (m_phase) throw NoriException("Medium: tried to register multiple PhaseFunction instances!");
            m_phase = static_cast<PhaseFunction *>(obj);
            break;
        case EEmitter:
            if (m_emitter) throw NoriException("Medium: tried to register multiple Emitter instances!");
            m_emitter = static_cast<Emitter *>(obj);
            break;
        default:
            throw NoriException("Medium::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
    }
}
void Medium::activate() {
    if (!m_phase) m_phase = static_cast<PhaseFunction *>(NoriObjectFactory::createInstance("isophase"));
}

class VacuumMedium : public Medium {
public:
    explicit VacuumMedium(const PropertyList &p) : Medium(p) {}
    bool getB200MediumData(nb_medium &out, FlatScene &) const override {
        memset(&out, 0, sizeof(out));
        out.type = NB_MEDIUM_VACUUM;
        return true;
    }
    std::string toString() const override { return "VacuumMedium[]"; }
};
NORI_REGISTER_CLASS(VacuumMedium, "vacuum")

class HomogeneousMedium : public Medium {           /* homogmedium.cpp:16-19, 48-60 */
public:
    explicit HomogeneousMedium(const PropertyList &p) : Medium(p), m_density(p.getFloat("density", 1.f)) {}
    bool getB200MediumData(nb_medium &out, FlatScene &) const override {
        memset(&out, 0, sizeof(out));
        out.type = NB_MEDIUM_HOMOGENEOUS;
        for (int i = 0; i < 3; ++i) out.mu_t[i] = m_density * m_sigma_a[i] + m_density * m_sigma_s[i];   /* mu_a + mu_s */
        m_phase->getB200PhaseData(out);
        return true;
    }
    std::string toString() const override { return "HomogeneousMedium[]"; }
private:
    float m_density;
};
NORI_REGISTER_CLASS(HomogeneousMedium, "homog")

/* <volume type="volume"> (src/textures/NvdbVolume.cpp, NvdbVolume.vdb.cpp): the "density" and
 * "temperature" float grids of an OpenVDB file. The reference converts them to NanoVDB through the
 * OpenVDB library and caches a .nvdb next to the file; here vdb.cpp reads the .vdb itself and the
 * grid goes to the GPU as a dense array (include/norib200.h, nb_medium). */
class NvdbVolume : public NoriObject {
public:
    explicit NvdbVolume(const PropertyList &p) : layout(p.getString("layout", "auto")), m_filename(resolvePath(p.getString("filename"))) {
        if (const char *env = getenv("NB_VOLUME_LAYOUT")) layout = env;
        if (layout != "auto" && layout != "dense" && layout != "bricks") throw NoriException("Volume: layout '%s' unknown (auto, dense, bricks)", layout);
    }
    EClassType getClassType() const override { return EVolume; }
    void activate() override {                          /* NvdbVolume::loadFromFile (NvdbVolume.vdb.cpp:9-40) */
        const size_t dot = m_filename.rfind('.');
        const std::string ext = dot == std::string::npos ? "" : m_filename.substr(dot);
        /* the reference's cache format (nvdb.cpp). Grid 0 of the file, whatever its name (NvdbVolume::loadNanoVdb, NvdbVolume.vdb.cpp:65:
         * readGrid(filename, 0, ..)); leaves kept as they are when the box is mostly empty */
        if (ext == ".nvdb") { density = nvdbReadFloatGrid(m_filename, "", layout != "dense"); return; }
        if (ext == ".nbv") { density = vdbReadNbv(m_filename); return; }     /* dense dump (vdb.h) */
        if (ext != ".vdb") throw NoriException("Volume: file extension %s unknown.", ext);
        density = vdbReadFloatGrid(m_filename, "density", layout != "dense");
    }
    std::string toString() const override { return "Volume[\n  filename = " + m_filename + ",\n]"; }
    VdbGrid density;
    /* how the grid goes to the GPU (include/norib200.h, nb_medium) — a property of this backend, not of the reference:
     * "dense", "bricks" (sparse), or "auto": bricks when they take less than half of what the dense box takes */
    std::string layout;
private:
    std::string m_filename;
};
NORI_REGISTER_CLASS(NvdbVolume, "volume")

/* The sparse device layout (include/norib200.h, NB_MEDIUM_HETEROGENEOUS_BRICKS): blocks of 8^3 voxels of the virtual array
 * (origin `org` in index space, `dims` voxels), each stored with a one-voxel apron — 10^3 floats — when anything in those 10^3
 * differs from the background; brick 0 is the background. A sparse source only visits the blocks its leaves can touch. */
static void buildBricks(const VdbGrid &g, const uint32_t dims[3], const int org[3], std::vector<uint32_t> &table, std::vector<float> &payload) {
    const int nb[3] = { (int) (dims[0] + 7) / 8, (int) (dims[1] + 7) / 8, (int) (dims[2] + 7) / 8 };
    const uint64_t entries = (uint64_t) nb[0] * nb[1] * nb[2];
    if (entries > (1ull << 30)) throw NoriException("HeterogeneousMedium: a brick table of %llu entries is too large", (unsigned long long) entries);
    table.assign((size_t) entries, 0u);
    payload.assign(NB_BRICK_VOXELS, g.background);
    std::vector<uint8_t> candidate;
    if (g.sparse()) {
        candidate.assign((size_t) entries, 0);
        for (const VdbLeaf &l : g.leaves) {
            int lo[3], hi[3];
            for (int a = 0; a < 3; ++a) {
                const int a0 = l.org[a] - org[a];                                   /* array coordinates of the leaf: a0 .. a0 + 7 */
                lo[a] = std::max(0, (a0 - 1) >> 3); hi[a] = std::min(nb[a] - 1, (a0 + 8) >> 3);
            }
            for (int i = lo[0]; i <= hi[0]; ++i)
                for (int j = lo[1]; j <= hi[1]; ++j)
                    for (int k = lo[2]; k <= hi[2]; ++k) candidate[((size_t) i * nb[1] + j) * nb[2] + k] = 1;
        }
    }
    float brick[NB_BRICK_VOXELS];
    for (int i = 0; i < nb[0]; ++i)
        for (int j = 0; j < nb[1]; ++j)
            for (int k = 0; k < nb[2]; ++k) {
                const size_t e = ((size_t) i * nb[1] + j) * nb[2] + k;
                if (!candidate.empty() && !candidate[e]) continue;
                bool any = false;
                int n = 0;
                for (int x = 0; x < NB_BRICK_EDGE; ++x)
                    for (int y = 0; y < NB_BRICK_EDGE; ++y)
                        for (int z = 0; z < NB_BRICK_EDGE; ++z) {
                            const float v = g.at(org[0] + i * 8 - 1 + x, org[1] + j * 8 - 1 + y, org[2] + k * 8 - 1 + z);
                            brick[n++] = v;
                            any = any || v != g.background;
                        }
                if (!any) continue;
                const uint64_t id = payload.size() / NB_BRICK_VOXELS;
                if (id >= 0x7f800000ull) throw NoriException("HeterogeneousMedium: too many bricks");
                table[e] = (uint32_t) id;
                payload.insert(payload.end(), brick, brick + NB_BRICK_VOXELS);
            }
}

class HeterogeneousMedium : public Medium {         /* heterogmedium.cpp:47-103 */
public:
    explicit HeterogeneousMedium(const PropertyList &p)
        : Medium(p), m_densityScale(p.getFloat("densityScale", 1.f)), m_temperatureScale(p.getFloat("temperatureScale", 0.f)) {}
    ~HeterogeneousMedium() override { delete m_volume; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != EVolume) {
            throw NoriException("HeterogeneousMedium::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        }
        if (m_volume) throw NoriException("There is already an volume defined!");
        m_volume = static_cast<NvdbVolume *>(obj);
    }
    void activate() override {
        if (!m_volume) throw NoriException("HeterogeneousMedium requires a nvdb volume.");
        Medium::activate();
    }
    bool getB200MediumData(nb_medium &out, FlatScene &flat) const override {
        memset(&out, 0, sizeof(out));
        const VdbGrid &g = m_volume->density;
        /* temperatureScale: the reference sums colorFromTemperature into a local it never returns (heterogmedium.cpp:152, :165-167) —
         * the property has no effect there, so it has none here. A volume light inside a heterogeneous medium: no kernel. */
        if (m_emitter) return false;
        if (g.active_voxels == 0 || !(g.value_max > 0.f)) return false;
        out.type = NB_MEDIUM_HETEROGENEOUS;                              /* or _BRICKS, below */
        for (int i = 0; i < 3; ++i) out.mu_t[i] = m_sigma_a[i] + m_sigma_s[i];
        m_phase->getB200PhaseData(out);
        out.density_scale = m_densityScale;
        out.max_density = g.value_max;
        /* virtual array = active box + border, read through getValue so stored inactive values survive */
        const int B = NB_GRID_BORDER;
        int org[3];
        for (int a = 0; a < 3; ++a) { out.grid_dims[a] = (uint32_t) (g.dims[a] + 2 * B); org[a] = g.bbox_min[a] - B; }
        const uint64_t vox = (uint64_t) out.grid_dims[0] * out.grid_dims[1] * out.grid_dims[2];
        if (flat.density.size() > 0xffffffffull) throw NoriException("HeterogeneousMedium: the density array is full");
        out.first_voxel = (uint32_t) flat.density.size();
        bool bricks = g.sparse() || m_volume->layout == "bricks";
        std::vector<uint32_t> table;
        std::vector<float> payload;
        if (bricks || (m_volume->layout == "auto" && vox >= (1ull << 24))) {
            buildBricks(g, out.grid_dims, org, table, payload);
            if (!bricks) bricks = (table.size() + payload.size()) * 2 < vox;
        }
        if (bricks) {
            out.type = NB_MEDIUM_HETEROGENEOUS_BRICKS;
            const size_t at = flat.density.size();
            flat.density.resize(at + table.size() + payload.size());
            memcpy(flat.density.data() + at, table.data(), table.size() * 4);
            memcpy(flat.density.data() + at + table.size(), payload.data(), payload.size() * 4);
        } else {
            if (vox > (1ull << 31)) throw NoriException("HeterogeneousMedium: a dense copy of the grid's %llu-voxel box does not fit; use layout=bricks", (unsigned long long) vox);
            flat.density.reserve(flat.density.size() + (size_t) vox);
            for (int i = 0; i < (int) out.grid_dims[0]; ++i)
                for (int j = 0; j < (int) out.grid_dims[1]; ++j)
                    for (int k = 0; k < (int) out.grid_dims[2]; ++k) flat.density.push_back(g.at(org[0] + i, org[1] + j, org[2] + k));
        }
        /* world -> index (inverse of the grid's affine map, in double), then index -> array coordinates */
        const double *M = g.index_to_world;
        const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], gg = M[8], h = M[9], k = M[10];
        const double det = a * (e * k - f * h) - b * (d * k - f * gg) + c * (d * h - e * gg);
        if (det == 0.0) throw NoriException("HeterogeneousMedium: the volume's transform is singular");
        const double inv[9] = { (e * k - f * h) / det, (c * h - b * k) / det, (b * f - c * e) / det,
                                (f * gg - d * k) / det, (a * k - c * gg) / det, (c * d - a * f) / det,
                                (d * h - e * gg) / det, (b * gg - a * h) / det, (a * e - b * d) / det };
        for (int r = 0; r < 3; ++r) {
            double t = 0.0;
            for (int cc = 0; cc < 3; ++cc) { out.world_to_grid[r * 4 + cc] = (float) inv[r * 3 + cc]; t -= inv[r * 3 + cc] * M[cc * 4 + 3]; }
            out.world_to_grid[r * 4 + 3] = (float) (t - org[r]);
        }
        return true;
    }
    std::string toString() const override { return "HeterogeneousMedium[...]"; }
private:
    NvdbVolume *m_volume = nullptr;
    float m_densityScale, m_temperatureScale;
};
NORI_REGISTER_CLASS(HeterogeneousMedium, "heterog")

Shape::~Shape() { delete m_bsdf; delete m_emitter; delete m_medium; delete m_normalMap; }

void Shape::addChild(NoriObject *obj) {
    switch (obj->getClassType()) {
        case EBSDF:
            if (m_bsdf) throw NoriException("Shape: tried to register multiple BSDF instances!");
            m_bsdf = static_cast<BSDF *>(obj);
            break;
        case EEmitter:
            if (m_emitter) throw NoriException("Shape: tried to register multiple Emitter instances!");
            m_emitter = static_cast<Emitter *>(obj);
            m_emitter->setShape(this);
            break;
        case EMedium:
            if (m_medium) throw NoriException("Shape: tried to register multiple medium instances!");
            m_medium = static_cast<Medium *>(obj);
            break;
        case ETexture:                                                   /* shape.cpp:139-148 */
            if (obj->getIdName() != "normal") throw NoriException("Shape does not have a texture with name: %s", obj->getIdName());
            if (m_normalMap) throw NoriException("There is already a normal map defined!");
            m_normalMap = static_cast<Texture *>(obj);
            break;
        default:
            throw NoriException("Shape::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
    }
}

void Shape::activate() {
    /* "If no material or medium was assigned, instantiate a diffuse BRDF" (shape.cpp:33-35) */
    if (!m_bsdf && !m_medium) m_bsdf = static_cast<BSDF *>(NoriObjectFactory::createInstance("diffuse"));
}

/* ---- Sphere (src/shapes/sphere.cpp:30-38) ---------------------------------------------------------------- */
class Sphere : public Shape {
public:
    explicit Sphere(const PropertyList &p) : m_position(p.getPoint3("center", Point3f())), m_radius(p.getFloat("radius", 1.f)) {
        m_bbox.expandBy(m_position - Vector3f(m_radius));
        m_bbox.expandBy(m_position + Vector3f(m_radius));
    }
    uint32_t getPrimitiveCount() const override { return 1; }
    BoundingBox3f getBoundingBox(uint32_t) const override { return m_bbox; }
    Point3f getCentroid(uint32_t) const override { return m_position; }
    float getVolume() const override { return (float) (4.f / 3.f * M_PI * std::pow(m_radius, 3)); }   /* sphere.h:45 */
    void getB200Shape(nb_shape &rec, FlatScene &) const override {
        rec.type = NB_SHAPE_SPHERE; rec.prim_count = 1; rec.radius = m_radius;
        for (int i = 0; i < 3; ++i) rec.center[i] = m_position[i];
    }
    std::string toString() const override { return NoriException::format("Sphere[radius=%f]", m_radius); }
private:
    Point3f m_position;
    float m_radius;
};
NORI_REGISTER_CLASS(Sphere, "sphere")

} // namespace nori
