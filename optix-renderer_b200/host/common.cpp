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

/* The reference inverts with Eigen 3.3.8's fp32 4x4 kernel (Transform ctor, transform.h:29 ->
 * Eigen/src/LU/arch/Inverse_SSE.h): the matrix is split into four 2x2 blocks
 *      M = | A C |      (blocks held column-wise: {m00, m10, m01, m11})
 *          | B D |
 * and the inverse assembled from adjugate products,
 *      det  = |A||D| + |B||C| - tr(A# B D# C)
 *      iA = A|D| - B (D# C)      iD = D|A| - C (A# B)
 *      iB = C|B| - D (A# B)#     iC = B|C| - A (D# C)#
 * scaled by +-1/det. Restated here four lanes at a time in the same operation order, so the
 * result (including the signs of its zeros) is the reference's bit for bit. */
namespace {
struct Quad { float v[4]; };
inline Quad lanes(const Quad &x, const Quad &y, int a, int b, int c, int d) { return Quad{{ x.v[a], x.v[b], y.v[c], y.v[d] }}; }
inline Quad operator*(const Quad &a, const Quad &b) { return Quad{{ a.v[0] * b.v[0], a.v[1] * b.v[1], a.v[2] * b.v[2], a.v[3] * b.v[3] }}; }
inline Quad operator-(const Quad &a, const Quad &b) { return Quad{{ a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2], a.v[3] - b.v[3] }}; }
inline Quad operator+(const Quad &a, const Quad &b) { return Quad{{ a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2], a.v[3] + b.v[3] }}; }
inline Quad splat(float f) { return Quad{{ f, f, f, f }}; }
/* A# * B for column-wise 2x2 blocks (A# = adjugate) */
inline Quad adj_mul(const Quad &A, const Quad &B) { return lanes(A, A, 3, 3, 0, 0) * B - lanes(A, A, 1, 1, 2, 2) * lanes(B, B, 2, 3, 0, 1); }
inline float det2(const Quad &A) { const Quad t = lanes(A, A, 3, 3, 1, 1) * A; return t.v[0] - t.v[2]; }
/* X * Y for column-wise blocks */
inline Quad blk_mul(const Quad &X, const Quad &Y) { return lanes(X, X, 0, 0, 2, 2) * lanes(Y, Y, 0, 1, 0, 1) + lanes(X, X, 1, 1, 3, 3) * lanes(Y, Y, 2, 3, 2, 3); }
/* X * (Y)# */
inline Quad mul_adj(const Quad &X, const Quad &Y) { return X * lanes(Y, Y, 3, 0, 3, 0) - lanes(X, X, 1, 0, 3, 2) * lanes(Y, Y, 2, 1, 2, 1); }
}

Matrix4f Matrix4f::inverse() const {
    auto col = [&](int k) { return Quad{{ m[0 * 4 + k], m[1 * 4 + k], m[2 * 4 + k], m[3 * 4 + k] }}; };
    const Quad c0 = col(0), c1 = col(1), c2 = col(2), c3 = col(3);
    const Quad A = lanes(c0, c1, 0, 1, 0, 1), B = lanes(c0, c1, 2, 3, 2, 3);
    const Quad C = lanes(c2, c3, 0, 1, 0, 1), D = lanes(c2, c3, 2, 3, 2, 3);
    const Quad AB = adj_mul(A, B), DC = adj_mul(D, C);
    const float dA = det2(A), dB = det2(B), dC = det2(C), dD = det2(D);
    Quad tr = lanes(DC, DC, 0, 2, 1, 3) * AB;
    Quad iD = blk_mul(C, AB);
    Quad iA = blk_mul(B, DC);
    const float t0 = tr.v[0] + tr.v[2], t1 = tr.v[1] + tr.v[3];
    const float trace = t0 + t1;
    const float d1 = dA * dD, d2 = dB * dC;
    iD = D * splat(dA) - iD;
    iA = A * splat(dD) - iA;
    const float det = (d1 + d2) - trace;
    const float rd = 1.0f / det;
    Quad iB = mul_adj(D, AB);
    Quad iC = mul_adj(A, DC);
    const Quad rdv = Quad{{ rd, -rd, -rd, rd }};
    iB = C * splat(dB) - iB;
    iC = B * splat(dC) - iC;
    iA = rdv * iA; iB = rdv * iB; iC = rdv * iC; iD = rdv * iD;
    const Quad r0 = lanes(iA, iB, 3, 1, 3, 1), r1 = lanes(iA, iB, 2, 0, 2, 0);
    const Quad r2 = lanes(iC, iD, 3, 1, 3, 1), r3 = lanes(iC, iD, 2, 0, 2, 0);
    Matrix4f out;
    for (int i = 0; i < 4; ++i) { out.m[i * 4 + 0] = r0.v[i]; out.m[i * 4 + 1] = r1.v[i]; out.m[i * 4 + 2] = r2.v[i]; out.m[i * 4 + 3] = r3.v[i]; }
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
