/* textures.cpp — Texture<Color3f> plugins (src/textures/consttexture.cpp, checkerboard.cpp,
 * PNGTexture.cpp) and the environment-map emitter built on them (src/emitters/environmentmap.cpp).
 * The textures evaluate on the host exactly as the reference's do — EnvMap::calculateProbs walks
 * them to build its luminance CDF — and flatten into nb_texture records for the kernels. */
#include "plugins.h"
#include "png.h"
#include <cstring>

namespace nori {

static Vector3f sphericalDirection(float theta, float phi) {             /* common.cpp:270-281 */
    const float st = std::sin(theta), ct = std::cos(theta), sp = std::sin(phi), cp = std::cos(phi);
    return Vector3f(st * cp, st * sp, ct);
}
static Point2f sphericalCoordinates(const Vector3f &v) {                 /* common.cpp:283-291 */
    Point2f r(std::acos(v.z()), std::atan2(v.y(), v.x()));
    if (r.v[1] < 0) r.v[1] += 2 * (float) M_PI;
    return r;
}
static float luminance(const Color3f &c) { return c[0] * 0.212671f + c[1] * 0.715160f + c[2] * 0.072169f; }

class ConstantTexture : public Texture {                                 /* consttexture.cpp:37, 62-65 */
public:
    explicit ConstantTexture(const PropertyList &p) : m_value(p.getColor("value", Color3f(0.f))) {}
    Color3f eval(const Point2f &) const override { return m_value; }
    int getB200Texture(FlatScene &flat) const override {
        nb_texture t; memset(&t, 0, sizeof(t));
        t.type = NB_TEXTURE_CONSTANT; t.width = t.height = 1;
        for (int i = 0; i < 3; ++i) t.value[i] = m_value[i];
        flat.textures.push_back(t);
        return (int) flat.textures.size() - 1;
    }
    std::string toString() const override { return NoriException::format("ConstantTexture[ %f %f %f ]", m_value[0], m_value[1], m_value[2]); }
private:
    Color3f m_value;
};
NORI_REGISTER_CLASS(ConstantTexture, "constant_color")

class Checkerboard : public Texture {                                    /* checkerboard.cpp:34-52, 81-88 */
public:
    explicit Checkerboard(const PropertyList &p)
        : m_delta(p.getPoint2("delta", Point2f(0, 0))), m_scale(p.getVector2("scale", Vector2f(1, 1))),
          m_value1(p.getColor("value1", Color3f(0.f))), m_value2(p.getColor("value2", Color3f(1.f))) {}
    Color3f eval(const Point2f &uv) const override {
        const float ox = uv.x() / m_scale.x() - m_delta.x(), oy = uv.y() / m_scale.y() - m_delta.y();
        const int x = int(ox) + (ox < 0.f), y = int(oy) + (oy < 0.f);
        return (x + y) % 2 == 0 ? m_value1 : m_value2;
    }
    int getB200Texture(FlatScene &flat) const override {
        nb_texture t; memset(&t, 0, sizeof(t));
        t.type = NB_TEXTURE_CHECKERBOARD; t.width = t.height = 1;
        for (int i = 0; i < 3; ++i) { t.value[i] = m_value1[i]; t.value2[i] = m_value2[i]; }
        t.scale[0] = m_scale.x(); t.scale[1] = m_scale.y(); t.offset[0] = m_delta.x(); t.offset[1] = m_delta.y();
        flat.textures.push_back(t);
        return (int) flat.textures.size() - 1;
    }
    std::string toString() const override { return "Checkerboard[]"; }
private:
    Point2f m_delta; Vector2f m_scale; Color3f m_value1, m_value2;
};
NORI_REGISTER_CLASS(Checkerboard, "checkerboard_color")

class PNGTexture : public Texture {                                      /* PNGTexture.cpp:23-165 */
public:
    explicit PNGTexture(const PropertyList &p) {
        m_filename = resolvePath(p.getString("filename"));
        m_sRgb = p.getBoolean("sRGB", p.getString("name", "") != "normal");
        m_scaleU = p.getFloat("scaleU", 1.f); m_scaleV = p.getFloat("scaleV", 1.f);
        m_euler = p.getVector3("eulerAngles", Vector3f(0.f)) * (float) M_PI / 180.f;
        m_spherical = p.getBoolean("sphericalTexture", false);
        m_offsetU = p.getFloat("offsetU", 0.f); m_offsetV = p.getFloat("offsetV", 0.f);
        m_intensity = p.getFloat("intensity", 1.f);
    }
    void activate() override {                                           /* loadFromFile, PNGTexture.cpp:68-123 */
        const size_t dot = m_filename.rfind('.');
        const std::string ext = dot == std::string::npos ? "" : m_filename.substr(dot);
        if (ext != ".png") { m_supported = false; return; }              /* .hdr (HDRLoader): no kernel, the scene stays on the CPU */
        std::vector<uint8_t> tmp;
        decodePng(m_filename, tmp, m_width, m_height);
        m_data.resize(tmp.size());
        if (m_sRgb) for (size_t i = 0; i < tmp.size(); ++i) m_data[i] = inverseGammaCorrect((float) tmp[i] / 255.f);
        else                                                             /* normal map, PNGTexture.cpp:88-99 as written: every third float */
            for (size_t i = 0; i < tmp.size(); ++i) {                    /* closes a triple that is normalized in place, across the RGBA stride */
                m_data[i] = (float) tmp[i] / 255 * 2 - 1;
                if ((i + 1) % 3 == 0) {
                    float *v = m_data.data() + i - 2;
                    const float z = v[0] * v[0] + (v[1] * v[1] + v[2] * v[2]);
                    if (z > 0.f) { const float n = std::sqrt(z); v[0] /= n; v[1] /= n; v[2] /= n; }
                }
            }
        /* the rotation PNGTexture::eval builds per call: Quaternion(Identity * AngleAxis(e.x, Z) * AngleAxis(e.y, X))
         * * AngleAxis(e.z, Z).toRotationMatrix() */
        float q[4] = { 1, 0, 0, 0 };                                     /* w, x, y, z */
        mulAxis(q, m_euler.x(), 2); mulAxis(q, m_euler.y(), 0);
        float Rq[9]; quatToMatrix(q, Rq);
        const float c = std::cos(m_euler.z()), s = std::sin(m_euler.z());
        const float Rz[9] = { c, -s, 0, s, c, 0, 0, 0, 1 };
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) m_rot[r * 3 + k] = Rq[r * 3] * Rz[k] + Rq[r * 3 + 1] * Rz[3 + k] + Rq[r * 3 + 2] * Rz[6 + k];
    }
    Color3f eval(const Point2f &_uv) const override {
        float u = _uv.x(), v = _uv.y();
        if (m_spherical) {
            const Vector3f wi = sphericalDirection(_uv.y() * (float) M_PI, _uv.x() * 2.f * (float) M_PI);
            const Vector3f w(m_rot[0] * wi[0] + m_rot[1] * wi[1] + m_rot[2] * wi[2], m_rot[3] * wi[0] + m_rot[4] * wi[1] + m_rot[5] * wi[2],
                             m_rot[6] * wi[0] + m_rot[7] * wi[1] + m_rot[8] * wi[2]);
            const Point2f c = sphericalCoordinates(w);
            u = c.y() / (2.f * (float) M_PI); v = c.x() / (float) M_PI;
        } else { u += m_offsetU; v += m_offsetV; }
        const unsigned w = static_cast<unsigned>(u * m_scaleU * (float) m_width);
        const unsigned h = m_height - static_cast<unsigned>(v * m_scaleV * (float) m_height);
        const unsigned index = (h * m_width + w) % (m_width * m_height);
        Color3f out(m_data[4 * index], m_data[4 * index + 1], m_data[4 * index + 2]);
        if (!m_sRgb) out = Color3f(out[0] * m_intensity, out[1] * m_intensity, out[2] * m_intensity + (1.f - m_intensity)).normalized();
        return out;
    }
    unsigned getWidth() const override { return m_width; }
    unsigned getHeight() const override { return m_height; }
    int getB200Texture(FlatScene &flat) const override {
        if (!m_supported || m_width == 0) return -1;
        nb_texture t; memset(&t, 0, sizeof(t));
        t.type = NB_TEXTURE_IMAGE; t.flags = (m_spherical ? NB_TEXTURE_SPHERICAL : 0u) | (m_sRgb ? 0u : NB_TEXTURE_NORMAL_MAP);
        t.intensity = m_intensity;
        t.width = m_width; t.height = m_height; t.first_texel = (uint32_t) (flat.texels.size() / 4);
        t.scale[0] = m_scaleU; t.scale[1] = m_scaleV; t.offset[0] = m_offsetU; t.offset[1] = m_offsetV;
        memcpy(t.rotation, m_rot, sizeof(m_rot));
        flat.texels.insert(flat.texels.end(), m_data.begin(), m_data.end());
        flat.textures.push_back(t);
        return (int) flat.textures.size() - 1;
    }
    std::string toString() const override { return "PNGTexture[\n  filename = " + m_filename + "\n]"; }
private:
    static float inverseGammaCorrect(float value) {                      /* PNGTexture.cpp:442-447 */
        if (value <= 0.04045f) return value * 1.f / 12.92f;
        return std::pow((value + 0.055f) * 1.f / 1.055f, 2.4f);
    }
    static void mulAxis(float q[4], float angle, int axis) {             /* q = q * AngleAxis(angle, unit axis) */
        const float ha = 0.5f * angle, w = std::cos(ha), sn = std::sin(ha);
        float b[4] = { w, 0, 0, 0 }; b[1 + axis] = sn;
        const float a[4] = { q[0], q[1], q[2], q[3] };
        q[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
        q[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
        q[2] = a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3];
        q[3] = a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1];
    }
    static void quatToMatrix(const float q[4], float R[9]) {             /* Eigen QuaternionBase::toRotationMatrix */
        const float w = q[0], x = q[1], y = q[2], z = q[3];
        const float tx = 2 * x, ty = 2 * y, tz = 2 * z, twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x,
                    tyy = ty * y, tyz = tz * y, tzz = tz * z;
        R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
        R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
        R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
    }
    std::string m_filename;
    bool m_sRgb = true, m_spherical = false, m_supported = true;
    float m_scaleU = 1, m_scaleV = 1, m_offsetU = 0, m_offsetV = 0, m_intensity = 1;
    Vector3f m_euler;
    unsigned m_width = 0, m_height = 0;
    std::vector<float> m_data;
    float m_rot[9] = { 1, 0, 0, 0, 1, 0, 0, 0, 1 };
};
NORI_REGISTER_CLASS(PNGTexture, "png_texture")

/* src/emitters/environmentmap.cpp:5-66, 146-161 */
class EnvMap : public Emitter {
public:
    explicit EnvMap(const PropertyList &p) { lightProb = p.getFloat("lightWeight", 1.f); m_radiance = p.getColor("radiance", Color3f(1.f)); }
    ~EnvMap() override { delete m_map; }
    bool isEnvMap() const override { return true; }
    void addChild(NoriObject *obj) override {
        if (obj->getClassType() != ETexture) throw NoriException("EnvMap::addChild(<%s>) is not supported!", classTypeName(obj->getClassType()));
        if (obj->getIdName() != "albedo") throw NoriException("The name of this texture does not match any field!");
        if (m_map) throw NoriException("There is already an albedo defined!");
        m_map = static_cast<Texture *>(obj);
    }
    void activate() override {
        if (!m_map) {                                                    /* constant 0.5 fallback, environmentmap.cpp:13-18 */
            PropertyList l; l.setColor("value", Color3f(0.5f));
            m_map = static_cast<Texture *>(NoriObjectFactory::createInstance("constant_color", l));
        }
    }
    bool getB200EmitterData(nb_emitter &o, FlatScene &flat) const override {
        memset(&o, 0, sizeof(o));
        o.type = NB_EMITTER_ENVMAP;
        o.shape = m_map->getB200Texture(flat);
        if (o.shape < 0) return false;
        for (int i = 0; i < 3; ++i) o.radiance[i] = m_radiance[i];
        /* calculateProbs (environmentmap.cpp:146-161): DiscretePDF over height x width entries, the
         * texture evaluated at (i / height, j / width); append + normalize as dpdf.h:48-56, 101-113 */
        const unsigned H = m_map->getHeight(), W = m_map->getWidth();
        const uint32_t offset = (uint32_t) flat.area_cdf.size();
        float norm = 0.f;
        if (!(W == 1 && H == 1)) {
            flat.area_cdf.reserve(flat.area_cdf.size() + (size_t) W * H + 1);
            flat.area_cdf.push_back(0.f);
            for (unsigned i = 0; i < H; ++i)
                for (unsigned j = 0; j < W; ++j)
                    flat.area_cdf.push_back(flat.area_cdf.back() + std::abs(luminance(m_map->eval(Point2f(i / (float) H, j / (float) W)))));
            const float sum = flat.area_cdf.back();
            if (sum > 0) {
                norm = 1.0f / sum;
                for (size_t k = offset + 1; k < flat.area_cdf.size(); ++k) flat.area_cdf[k] *= norm;
                flat.area_cdf.back() = 1.0f;
            }
        }
        memcpy(&o.position[0], &offset, 4);
        o.position[1] = norm;
        return true;
    }
    std::string toString() const override { return "PngEnvMap[\n  texture = " + (m_map ? m_map->toString() : std::string("null")) + "\n]"; }
private:
    Texture *m_map = nullptr;
};
NORI_REGISTER_CLASS(EnvMap, "envmap")

}
