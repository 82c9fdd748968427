/* common.cpp — math, string conversion, object factory (mirror of src/utils/common.cpp:60-160,
 * src/utils/transform.cpp, src/utils/object.cpp). */
#include "nori.h"
#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace nori {

Matrix4f Matrix4f::operator*(const Matrix4f &o) const {
    Matrix4f r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            float acc = m[i * 4 + 0] * o.m[0 * 4 + j];
            acc = acc + m[i * 4 + 1] * o.m[1 * 4 + j];
            acc = acc + m[i * 4 + 2] * o.m[2 * 4 + j];
            acc = acc + m[i * 4 + 3] * o.m[3 * 4 + j];
            r.m[i * 4 + j] = acc;
        }
    return r;
}

Matrix4f Matrix4f::inverse() const {
    /* Gauss-Jordan with partial pivoting in double; the reference uses Eigen's 4x4 inverse in
     * fp32 — results agree to ~1 ulp, which only shifts camera rays / normals within noise. */
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { a[i][j] = m[i * 4 + j]; a[i][4 + j] = (i == j) ? 1.0 : 0.0; }
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
        if (a[piv][c] == 0.0) throw NoriException("Matrix4f::inverse(): singular matrix");
        if (piv != c) for (int j = 0; j < 8; ++j) std::swap(a[piv][j], a[c][j]);
        const double d = a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] /= d;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            const double f = a[r][c];
            if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    Matrix4f out;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out.m[i * 4 + j] = (float) a[i][4 + j];
    return out;
}

Point3f Transform::point(const Point3f &p) const {
    float r[4];
    for (int i = 0; i < 4; ++i) {
        /* Eigen's 4x4 * 4x1 product accumulates column by column */
        float acc = m.m[i * 4 + 0] * p[0];
        acc = acc + m.m[i * 4 + 1] * p[1];
        acc = acc + m.m[i * 4 + 2] * p[2];
        acc = acc + m.m[i * 4 + 3] * 1.0f;
        r[i] = acc;
    }
    return Point3f(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
}

Vector3f Transform::vector(const Vector3f &v) const {
    Vector3f r;
    for (int i = 0; i < 3; ++i) {
        float acc = m.m[i * 4 + 0] * v[0];
        acc = acc + m.m[i * 4 + 1] * v[1];
        acc = acc + m.m[i * 4 + 2] * v[2];
        r[i] = acc;
    }
    return r;
}

Normal3f Transform::normal(const Normal3f &n) const {
    Vector3f r;
    for (int i = 0; i < 3; ++i) {          /* inverse^T * n */
        float acc = inv.m[0 * 4 + i] * n[0];
        acc = acc + inv.m[1 * 4 + i] * n[1];
        acc = acc + inv.m[2 * 4 + i] * n[2];
        r[i] = acc;
    }
    return r;
}

std::vector<std::string> tokenize(const std::string &s, const std::string &delim, bool includeEmpty) {
    std::vector<std::string> out;
    size_t last = 0, pos = s.find_first_of(delim, last);
    while (last != std::string::npos) {
        if (pos != last || includeEmpty) out.push_back(s.substr(last, pos - last));
        last = pos;
        if (last != std::string::npos) { last += 1; pos = s.find_first_of(delim, last); }
    }
    return out;
}

std::string toLower(const std::string &s) {
    std::string r(s);
    std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return (char) std::tolower(c); });
    return r;
}

bool toBool(const std::string &s) {
    const std::string v = toLower(s);
    if (v == "false") return false;
    if (v == "true") return true;
    throw NoriException("Could not parse boolean value \"%s\"", s);
}

int toInt(const std::string &s) {
    char *end = nullptr;
    int r = (int) strtol(s.c_str(), &end, 10);
    if (*end != '\0') throw NoriException("Could not parse integer value \"%s\"", s);
    return r;
}

float toFloat(const std::string &s) {
    char *end = nullptr;
    float r = strtof(s.c_str(), &end);
    if (*end != '\0') throw NoriException("Could not parse floating point value \"%s\"", s);
    return r;
}

Vector3f toVector3f(const std::string &s) {
    std::vector<std::string> t = tokenize(s);
    if (t.size() != 3) throw NoriException("Expected 3 values");
    return Vector3f(toFloat(t[0]), toFloat(t[1]), toFloat(t[2]));
}

Vector2f toVector2f(const std::string &s) {
    std::vector<std::string> t = tokenize(s);
    if (t.size() != 2) throw NoriException("Expected 2 values");
    return Vector2f(toFloat(t[0]), toFloat(t[1]));
}

const char *NoriObject::classTypeName(EClassType t) {
    switch (t) {
        case EScene: return "scene"; case EShape: return "shape"; case ETexture: return "texture";
        case EVolume: return "volume"; case EBSDF: return "bsdf"; case EPhaseFunction: return "phase";
        case EEmitter: return "emitter"; case EMedium: return "medium"; case ECamera: return "camera";
        case EIntegrator: return "integrator"; case ESampler: return "sampler"; case EPixelSampler: return "pxsampler";
        case EDenoiser: return "denoiser"; case ETest: return "test"; case EReconstructionFilter: return "rfilter";
        case ERenderer: return "renderer"; default: return "<unknown>";
    }
}

static std::map<std::string, NoriObjectFactory::Constructor> &registry() {
    static std::map<std::string, NoriObjectFactory::Constructor> r;
    return r;
}

void NoriObjectFactory::registerClass(const std::string &name, const Constructor &c) { registry()[name] = c; }
bool NoriObjectFactory::has(const std::string &name) { return registry().count(name) != 0; }

NoriObject *NoriObjectFactory::createInstance(const std::string &name, const PropertyList &props) {
    auto it = registry().find(name);
    if (it == registry().end())
        throw NoriException("A constructor for class \"%s\" could not be found!", name);
    return it->second(props);
}

} // namespace nori
