/* nori.h — host-side mirror of the reference's scene/plugin surface, written from scratch.
 *
 * Same names and meaning as the reference so scene XML written for it loads unchanged:
 * NoriObject / EClassType / NoriObjectFactory / NORI_REGISTER_CLASS (include/nori/object.h:38-291),
 * PropertyList (include/nori/proplist.h:32-161), NoriException (include/nori/common.h:153-158),
 * loadFromXML (src/utils/parser.cpp:28-378). The plugins here hold parameters and implement the
 * *flatten hooks* that feed the C ABI (the role of getOptixMaterialData & co., bsdf.h:122,
 * emitter.h:159, camera.h:86, shape.h:194-201); the light-transport math itself lives only in the
 * CUDA kernels — there is no CPU renderer in this library.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/norib200.h"

namespace nori {

/* ---- errors ------------------------------------------------------------------------------------ */
class NoriException : public std::runtime_error {
public:
    explicit NoriException(const std::string &s) : std::runtime_error(s) {}
    template <typename... Args> NoriException(const char *fmt, Args... args) : std::runtime_error(format(fmt, args...)) {}
    template <typename... Args> static std::string format(const char *fmt, Args... args) {
        char buf[2048];
        snprintf(buf, sizeof(buf), fmt, cstr(args)...);
        return buf;
    }
private:
    template <typename T> static T cstr(T v) { return v; }
    static const char *cstr(const std::string &s) { return s.c_str(); }
};

/* ---- small linear algebra (fp32, operation order chosen to follow Eigen where it matters) ------ */
struct Vector3f {
    float v[3];
    Vector3f() : v{0, 0, 0} {}
    Vector3f(float x, float y, float z) : v{x, y, z} {}
    explicit Vector3f(float s) : v{s, s, s} {}
    float &operator[](int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
    float x() const { return v[0]; }
    float y() const { return v[1]; }
    float z() const { return v[2]; }
    Vector3f operator+(const Vector3f &o) const { return Vector3f(v[0] + o.v[0], v[1] + o.v[1], v[2] + o.v[2]); }
    Vector3f operator-(const Vector3f &o) const { return Vector3f(v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]); }
    Vector3f operator*(float s) const { return Vector3f(v[0] * s, v[1] * s, v[2] * s); }
    Vector3f operator/(float s) const { return Vector3f(v[0] / s, v[1] / s, v[2] / s); }
    Vector3f operator-() const { return Vector3f(-v[0], -v[1], -v[2]); }
    float dot(const Vector3f &o) const { return v[0] * o.v[0] + (v[1] * o.v[1] + v[2] * o.v[2]); }   /* Eigen redux order */
    Vector3f cross(const Vector3f &o) const {
        return Vector3f(v[1] * o.v[2] - v[2] * o.v[1], v[2] * o.v[0] - v[0] * o.v[2], v[0] * o.v[1] - v[1] * o.v[0]);
    }
    float squaredNorm() const { return dot(*this); }
    float norm() const { return std::sqrt(squaredNorm()); }
    Vector3f normalized() const { float z = squaredNorm(); return z > 0.f ? *this / std::sqrt(z) : *this; }
    float maxCoeff() const { return std::max(v[0], std::max(v[1], v[2])); }
};
typedef Vector3f Point3f;
typedef Vector3f Normal3f;
typedef Vector3f Color3f;
struct Vector2f { float v[2]; Vector2f() : v{0, 0} {} Vector2f(float x, float y) : v{x, y} {} float x() const { return v[0]; } float y() const { return v[1]; } };
typedef Vector2f Point2f;
struct Vector2i { int v[2]; Vector2i() : v{0, 0} {} Vector2i(int x, int y) : v{x, y} {} int x() const { return v[0]; } int y() const { return v[1]; } };

/* Row-major 4x4 (element (r,c) = m[r*4+c]) */
struct Matrix4f {
    float m[16];
    static Matrix4f identity() { Matrix4f r; for (int i = 0; i < 16; ++i) r.m[i] = (i % 5 == 0) ? 1.f : 0.f; return r; }
    float &operator()(int r, int c) { return m[r * 4 + c]; }
    float operator()(int r, int c) const { return m[r * 4 + c]; }
    Matrix4f operator*(const Matrix4f &o) const;
    Matrix4f inverse() const;     /* Eigen's fp32 block-adjugate scheme, operation for operation */
};

/* Transform (include/nori/transform.h:20-100): matrix + inverse */
struct Transform {
    Matrix4f m, inv;
    Transform() : m(Matrix4f::identity()), inv(Matrix4f::identity()) {}
    explicit Transform(const Matrix4f &t) : m(t), inv(t.inverse()) {}
    Point3f point(const Point3f &p) const;        /* transform.h:83-86 */
    Vector3f vector(const Vector3f &v) const;     /* transform.h:73-75 */
    Normal3f normal(const Normal3f &n) const;     /* transform.h:78-80: inverse transpose */
    const Matrix4f &getMatrix() const { return m; }
};

struct BoundingBox3f {                            /* include/nori/bbox.h */
    Point3f min, max;
    BoundingBox3f() { reset(); }
    explicit BoundingBox3f(const Point3f &p) : min(p), max(p) {}
    BoundingBox3f(const Point3f &a, const Point3f &b) : min(a), max(b) {}
    void reset() { min = Point3f(INFINITY); max = Point3f(-INFINITY); }
    void expandBy(const Point3f &p) { for (int i = 0; i < 3; ++i) { min[i] = std::min(min[i], p[i]); max[i] = std::max(max[i], p[i]); } }
    void expandBy(const BoundingBox3f &b) { for (int i = 0; i < 3; ++i) { min[i] = std::min(min[i], b.min[i]); max[i] = std::max(max[i], b.max[i]); } }
    static BoundingBox3f merge(const BoundingBox3f &a, const BoundingBox3f &b) { BoundingBox3f r(a); r.expandBy(b); return r; }
    float getSurfaceArea() const {               /* bbox.h:87-100, same accumulation order */
        Vector3f d = max - min;
        float result = 0.0f;
        for (int i = 0; i < 3; ++i) { float term = 1.0f; for (int j = 0; j < 3; ++j) { if (i == j) continue; term *= d[j]; } result += term; }
        return 2.0f * result;
    }
    int getLargestAxis() const {                 /* bbox.h:308-317 */
        Vector3f e = max - min;
        if (e[0] >= e[1] && e[0] >= e[2]) return 0;
        if (e[1] >= e[0] && e[1] >= e[2]) return 1;
        return 2;
    }
};

/* ---- string helpers (src/utils/common.cpp:60-160) ------------------------------------------------ */
std::vector<std::string> tokenize(const std::string &s, const std::string &delim = ", ", bool includeEmpty = false);
std::string toLower(const std::string &s);
bool toBool(const std::string &s);
int toInt(const std::string &s);
float toFloat(const std::string &s);
Vector3f toVector3f(const std::string &s);
Vector2f toVector2f(const std::string &s);
inline float degToRad(float v) { return v * (float) (M_PI / 180.0); }

/* ---- PropertyList (include/nori/proplist.h) ------------------------------------------------------- */
class PropertyList {
public:
    void setBoolean(const std::string &n, bool v) { set(n, EBool).b = v; }
    void setInteger(const std::string &n, int v) { set(n, EInt).i = v; }
    void setFloat(const std::string &n, float v) { set(n, EFloat).f = v; }
    void setString(const std::string &n, const std::string &v) { set(n, EString).s = v; }
    void setColor(const std::string &n, const Color3f &v) { set(n, EColor).v3 = v; }
    void setPoint3(const std::string &n, const Point3f &v) { set(n, EPoint3).v3 = v; }
    void setVector3(const std::string &n, const Vector3f &v) { set(n, EVector3).v3 = v; }
    void setPoint2(const std::string &n, const Point2f &v) { set(n, EPoint2).v2 = v; }
    void setVector2(const std::string &n, const Vector2f &v) { set(n, EVector2).v2 = v; }
    void setTransform(const std::string &n, const Transform &v) { set(n, ETransform).t = v; }
    bool has(const std::string &n) const { return m_props.count(n) != 0; }
    bool getBoolean(const std::string &n) const { return get(n, EBool, "boolean").b; }
    bool getBoolean(const std::string &n, bool d) const { return has(n) ? getBoolean(n) : d; }
    int getInteger(const std::string &n) const { return get(n, EInt, "integer").i; }
    int getInteger(const std::string &n, int d) const { return has(n) ? getInteger(n) : d; }
    float getFloat(const std::string &n) const { return get(n, EFloat, "float").f; }
    float getFloat(const std::string &n, float d) const { return has(n) ? getFloat(n) : d; }
    std::string getString(const std::string &n) const { return get(n, EString, "string").s; }
    std::string getString(const std::string &n, const std::string &d) const { return has(n) ? getString(n) : d; }
    Color3f getColor(const std::string &n) const { return get(n, EColor, "color").v3; }
    Color3f getColor(const std::string &n, const Color3f &d) const { return has(n) ? getColor(n) : d; }
    Point3f getPoint3(const std::string &n) const { return get(n, EPoint3, "point").v3; }
    Point3f getPoint3(const std::string &n, const Point3f &d) const { return has(n) ? getPoint3(n) : d; }
    Vector3f getVector3(const std::string &n) const { return get(n, EVector3, "vector").v3; }
    Vector3f getVector3(const std::string &n, const Vector3f &d) const { return has(n) ? getVector3(n) : d; }
    Vector2f getVector2(const std::string &n, const Vector2f &d) const { return has(n) ? get(n, EVector2, "vector2").v2 : d; }
    Point2f getPoint2(const std::string &n, const Point2f &d) const { return has(n) ? get(n, EPoint2, "point2").v2 : d; }
    Transform getTransform(const std::string &n) const { return get(n, ETransform, "transform").t; }
    Transform getTransform(const std::string &n, const Transform &d) const { return has(n) ? getTransform(n) : d; }
private:
    enum EType { EBool, EInt, EFloat, EString, EColor, EPoint3, EVector3, EPoint2, EVector2, ETransform };
    struct Property { EType type; bool b = false; int i = 0; float f = 0; std::string s; Vector3f v3; Vector2f v2; Transform t; };
    Property &set(const std::string &n, EType t) {
        if (m_props.count(n)) fprintf(stderr, "Property \"%s\" was specified multiple times!\n", n.c_str());
        Property &p = m_props[n]; p.type = t; return p;
    }
    const Property &get(const std::string &n, EType t, const char *tn) const {
        auto it = m_props.find(n);
        if (it == m_props.end()) throw NoriException("Property '%s' is missing!", n);
        if (it->second.type != t) throw NoriException("Property '%s' has the wrong type! (expected <%s>)!", n, tn);
        return it->second;
    }
    std::map<std::string, Property> m_props;
};

/* ---- NoriObject + factory (include/nori/object.h) -------------------------------------------------- */
class NoriObject {
public:
    enum EClassType { EScene = 0, EShape, ETexture, EVolume, EBSDF, EPhaseFunction, EEmitter, EMedium, ECamera,
                      EIntegrator, ESampler, EPixelSampler, EDenoiser, ETest, EReconstructionFilter, ERenderer,
                      EClassTypeCount };
    virtual ~NoriObject() {}
    virtual EClassType getClassType() const = 0;
    virtual void addChild(NoriObject *child) {
        throw NoriException("NoriObject::addChild() is not implemented for objects of type '%s'!", classTypeName(getClassType()));
    }
    virtual void setParent(NoriObject *) {}
    virtual void activate() {}
    void setIdName(const std::string &name) { m_idname = name; }        /* object.h:182-183: the XML `name` attribute */
    const std::string &getIdName() const { return m_idname; }
    virtual std::string toString() const = 0;
    static const char *classTypeName(EClassType t);
private:
    std::string m_idname;
};

class NoriObjectFactory {
public:
    typedef std::function<NoriObject *(const PropertyList &)> Constructor;
    static void registerClass(const std::string &name, const Constructor &c);
    static NoriObject *createInstance(const std::string &name, const PropertyList &props = PropertyList());
    static bool has(const std::string &name);
};

#define NORI_REGISTER_CLASS(cls, name) \
    static struct cls##_reg_ { cls##_reg_() { ::nori::NoriObjectFactory::registerClass(name, \
        [](const ::nori::PropertyList &l) -> ::nori::NoriObject * { return new cls(l); }); } } cls##_reg_instance_;

/* src/utils/parser.cpp:28 */
NoriObject *loadFromXML(const std::string &filename);

} // namespace nori
