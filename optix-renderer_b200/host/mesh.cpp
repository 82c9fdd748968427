/* mesh.cpp — Mesh (src/shapes/mesh.cpp) and the Wavefront OBJ loader (src/shapes/obj.cpp:77-238):
 * toWorld baked into positions/normals while reading, (v, vt, vn) triples deduplicated in order
 * of first use, quads split 0-1-2 / 3-0-2, per-vertex tangents/bitangents accumulated per
 * triangle; area CDF as DiscretePDF builds it (include/nori/dpdf.h:40-113). */
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <cstdio>
#include <thread>
#include <functional>
#include "plugins.h"
#include <fstream>
#include <map>
#include <sstream>
#include <tuple>

namespace nori {

BoundingBox3f Mesh::getBoundingBox(uint32_t index) const {         /* mesh.cpp:198-204 */
    BoundingBox3f r(vertex(m_F[3 * index]));
    r.expandBy(vertex(m_F[3 * index + 1]));
    r.expandBy(vertex(m_F[3 * index + 2]));
    return r;
}

Point3f Mesh::getCentroid(uint32_t index) const {                   /* mesh.cpp:206-212 */
    return (vertex(m_F[3 * index]) + vertex(m_F[3 * index + 1]) + vertex(m_F[3 * index + 2])) * (1.0f / 3.0f);
}

float Mesh::surfaceArea(uint32_t index) const {                     /* mesh.cpp:92-99 */
    const Point3f p0 = vertex(m_F[3 * index]), p1 = vertex(m_F[3 * index + 1]), p2 = vertex(m_F[3 * index + 2]);
    return 0.5f * (p1 - p0).cross(p2 - p0).norm();
}

void Mesh::activate() {
    Shape::activate();
    const uint32_t n = getPrimitiveCount();
    m_cdf.clear();
    m_cdf.reserve(n + 1);
    m_cdf.push_back(0.0f);
    for (uint32_t i = 0; i < n; ++i) m_cdf.push_back(m_cdf.back() + surfaceArea(i));
    const float sum = m_cdf.back();                                 /* DiscretePDF::normalize */
    if (sum > 0) {
        m_cdfNormalization = 1.0f / sum;
        for (size_t i = 1; i < m_cdf.size(); ++i) m_cdf[i] *= m_cdfNormalization;
        m_cdf[m_cdf.size() - 1] = 1.0f;
    } else m_cdfNormalization = 0.0f;
}

void Mesh::getB200Shape(nb_shape &rec, FlatScene &flat) const {
    const uint32_t nv = (uint32_t) (m_V.size() / 3);
    rec.type = NB_SHAPE_MESH;
    rec.prim_count = getPrimitiveCount();
    rec.vertex_offset = (uint32_t) (flat.positions.size() / 3);
    rec.vertex_count = nv;
    rec.face_offset = (uint32_t) (flat.faces.size() / 3);
    if (!m_N.empty()) rec.flags |= NB_SHAPE_HAS_NORMALS;
    if (!m_UV.empty()) rec.flags |= NB_SHAPE_HAS_UVS;
    flat.positions.insert(flat.positions.end(), m_V.begin(), m_V.end());
    /* optional attributes are parallel to positions: pad with zeros where a mesh has none */
    auto put = [&](std::vector<float> &dst, const std::vector<float> &src, bool wanted, size_t per) {
        if (!wanted) return;
        dst.resize((size_t) rec.vertex_offset * per, 0.f);
        if (!src.empty()) dst.insert(dst.end(), src.begin(), src.end());
        else dst.resize(dst.size() + (size_t) nv * per, 0.f);
    };
    put(flat.normals, m_N, flat.any_normals, 3);
    put(flat.uvs, m_UV, flat.any_uvs, 2);
    put(flat.tangents, m_T, flat.any_tangents, 3);
    put(flat.bitangents, m_BT, flat.any_tangents, 3);
    flat.faces.insert(flat.faces.end(), m_F.begin(), m_F.end());
    rec.cdf_offset = (uint32_t) flat.area_cdf.size();
    flat.area_cdf.insert(flat.area_cdf.end(), m_cdf.begin(), m_cdf.end());
    rec.pdf_norm = m_cdfNormalization;
}

std::string Mesh::toString() const {
    return NoriException::format("Mesh[name=\"%s\", vertexCount=%d, triangleCount=%d]", m_name, (int) (m_V.size() / 3), (int) (m_F.size() / 3));
}

class WavefrontOBJ : public Mesh {
public:
    explicit WavefrontOBJ(const PropertyList &p) {
        m_filename = resolvePath(p.getString("filename"));
        m_trafo = p.getTransform("toWorld", Transform());
        load();
    }
private:
    struct OBJVertex { uint32_t p = (uint32_t) -1, n = (uint32_t) -1, uv = (uint32_t) -1; };
    static uint32_t toUInt(const std::string &s) {
        char *end = nullptr;
        unsigned int r = (unsigned int) strtoul(s.c_str(), &end, 10);
        if (*end != '\0') throw NoriException("Could not parse integer value \"%s\"", s);
        return r;
    }
    static OBJVertex parseVertex(const std::string &s) {            /* obj.cpp:312-326 */
        std::vector<std::string> t = tokenize(s, "/", true);
        if (t.size() < 1 || t.size() > 3) throw NoriException("Invalid vertex data: \"%s\"", s);
        OBJVertex v;
        v.p = toUInt(t[0]);
        if (t.size() >= 2 && !t[1].empty()) v.uv = toUInt(t[1]);
        if (t.size() >= 3 && !t[2].empty()) v.n = toUInt(t[2]);
        return v;
    }
    /* The reference reads the file line by line through std::istringstream (obj.cpp:60-140): 4.6 s for the 49 MB, 1.3 M-triangle
     * mesh of C3 — more than a hundred times the render it feeds once the tree is built on the device. Same grammar, same
     * values (strtof rounds like operator>>(float)), same vertex numbering (first appearance of a p/uv/n triple), read from
     * one buffer with a hash table in place of the ordered map. A number that does not parse reads as 0 and ends the line's
     * numbers, as a failed stream extraction does. */
    static const char *skipSpace(const char *p, const char *end) { while (p < end && (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\v' || *p == '\f')) ++p; return p; }
    static const char *tokenEnd(const char *p, const char *end) { while (p < end && !(*p == ' ' || *p == '\t' || *p == '\r' || *p == '\v' || *p == '\f')) ++p; return p; }
    static int readFloats(const char *p, const char *end, float *out, int n) {
        int got = 0;
        for (; got < n; ++got) {
            p = skipSpace(p, end);
            if (p >= end) break;
            char *stop = nullptr;
            const float v = strtof(p, &stop);             /* the buffer ends with a NUL; a line ends with one too (see load) */
            if (stop == p) break;
            out[got] = v;
            p = stop;
        }
        return got;
    }
    static OBJVertex parseVertexFast(const char *p, const char *end) {       /* obj.cpp:312-326: p[/uv[/n]] */
        OBJVertex v;
        uint32_t *field[3] = { &v.p, &v.uv, &v.n };
        int f = 0;
        const char *tok = p;
        for (;; ++f) {
            if (f > 2) throw NoriException("Invalid vertex data: \"%s\"", std::string(tok, end));
            const char *q = p;
            uint64_t r = 0;
            while (q < end && *q >= '0' && *q <= '9') { r = r * 10 + (uint64_t) (*q - '0'); ++q; }
            if (q < end && *q != '/') throw NoriException("Could not parse integer value \"%s\"", std::string(p, tokenEndSlash(p, end)));
            if (q > p) *field[f] = (uint32_t) r;
            else if (f == 0) throw NoriException("Could not parse integer value \"\"");
            if (q >= end) break;
            p = q + 1;
        }
        return v;
    }
    static const char *tokenEndSlash(const char *p, const char *end) { while (p < end && *p != '/') ++p; return p; }
    void load() {
        const auto tStart = std::chrono::steady_clock::now();
        auto lap = [&](const char *what) { if (getenv("NB_HOST_TIMING")) fprintf(stderr, "obj %s: %.3f s\n", what, std::chrono::duration<double>(std::chrono::steady_clock::now() - tStart).count()); };
        FILE *file = fopen(m_filename.c_str(), "rb");
        if (!file) throw NoriException("Unable to open OBJ file \"%s\"!", m_filename);
        std::vector<char> buf;
        {
            fseek(file, 0, SEEK_END);
            const long size = ftell(file);
            fseek(file, 0, SEEK_SET);
            buf.resize((size_t) std::max(0L, size) + 1);
            const size_t got = fread(buf.data(), 1, buf.size() - 1, file);
            fclose(file);
            buf.resize(got + 1);
            buf[got] = '\0';
        }
        std::vector<Vector3f> positions, normals;
        std::vector<Vector2f> texcoords;
        std::vector<uint32_t> indices;
        std::vector<OBJVertex> vertices, corners;
        /* open addressing over (p, uv, n): the slot holds the vertex number + 1 */
        std::vector<uint32_t> table(1u << 16, 0u);
        auto hashOf = [](const OBJVertex &v) { uint64_t h = v.p * 0x9E3779B97F4A7C15ull; h ^= (h >> 29) + v.uv * 0xBF58476D1CE4E5B9ull; h ^= (h >> 31) + v.n * 0x94D049BB133111EBull; return h ^ (h >> 32); };
        std::vector<uint32_t> hashed;                                  /* the vertices the table holds (the others are found by position) */
        auto lookup = [&](const OBJVertex &v) -> uint32_t {
            if ((hashed.size() + 1) * 2 > table.size()) {             /* grow and re-insert */
                std::vector<uint32_t> bigger(table.size() * 4, 0u);
                for (uint32_t id : hashed) {
                    size_t at = hashOf(vertices[id]) & (bigger.size() - 1);
                    while (bigger[at]) at = (at + 1) & (bigger.size() - 1);
                    bigger[at] = id + 1;
                }
                table.swap(bigger);
            }
            size_t at = hashOf(v) & (table.size() - 1);
            for (;; at = (at + 1) & (table.size() - 1)) {
                const uint32_t id = table[at];
                if (!id) {
                    table[at] = (uint32_t) vertices.size() + 1;
                    hashed.push_back((uint32_t) vertices.size());
                    vertices.push_back(v);
                    return (uint32_t) vertices.size() - 1;
                }
                const OBJVertex &o = vertices[id - 1];
                if (o.p == v.p && o.uv == v.uv && o.n == v.n) return id - 1;
            }
        };
        /* the text is cut at line ends into one piece per host thread; the pieces are read side by side and joined in file
         * order, so the arrays are the ones a single pass gives. Corners are numbered afterwards, in file order. */
        struct Piece {
            char *begin = nullptr, *end = nullptr;
            std::vector<Vector3f> positions, normals;
            std::vector<Vector2f> texcoords;
            std::vector<OBJVertex> corners;
            BoundingBox3f bbox;
            std::string error;
        };
        char *const fileBegin = buf.data(), *const fileEnd = buf.data() + buf.size() - 1;
        const size_t bytes = (size_t) (fileEnd - fileBegin);
        size_t nPieces = std::max<size_t>(1, std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), bytes >> 21));
        if (const char *env = getenv("NB_OBJ_THREADS")) nPieces = (size_t) std::max(1, atoi(env));
        std::vector<Piece> pieces(nPieces);
        {
            char *at = fileBegin;
            for (size_t k = 0; k < nPieces; ++k) {
                pieces[k].begin = at;
                char *stop = k + 1 == nPieces ? fileEnd : fileBegin + bytes * (k + 1) / nPieces;
                if (stop < at) stop = at;
                if (stop < fileEnd) { char *nl = (char *) memchr(stop, '\n', (size_t) (fileEnd - stop)); stop = nl ? nl + 1 : fileEnd; }
                pieces[k].end = stop;
                at = stop;
            }
        }
        auto readPiece = [this](Piece &pc) {
            try {
                char *cur = pc.begin;
                while (cur < pc.end) {
                    char *eol = (char *) memchr(cur, '\n', (size_t) (pc.end - cur));
                    if (!eol) eol = pc.end;
                    *eol = '\0';                                      /* strtof stops at the end of the line (the buffer has one byte to spare) */
                    const char *p = skipSpace(cur, eol), *pe = tokenEnd(p, eol);
                    const size_t plen = (size_t) (pe - p);
                    if (plen == 1 && p[0] == 'v') {
                        float v[3] = { 0.f, 0.f, 0.f };
                        readFloats(pe, eol, v, 3);
                        const Point3f pt = m_trafo.point(Point3f(v[0], v[1], v[2]));
                        pc.bbox.expandBy(pt);
                        pc.positions.push_back(pt);
                    } else if (plen == 2 && p[0] == 'v' && p[1] == 't') {
                        float v[2] = { 0.f, 0.f };
                        readFloats(pe, eol, v, 2);
                        pc.texcoords.push_back(Vector2f(v[0], v[1]));
                    } else if (plen == 2 && p[0] == 'v' && p[1] == 'n') {
                        float v[3] = { 0.f, 0.f, 0.f };
                        readFloats(pe, eol, v, 3);
                        pc.normals.push_back(m_trafo.normal(Normal3f(v[0], v[1], v[2])).normalized());
                    } else if (plen == 1 && p[0] == 'f') {
                        OBJVertex verts[6];
                        int nTokens = 0;
                        const char *q = pe;
                        for (; nTokens < 4; ++nTokens) {
                            q = skipSpace(q, eol);
                            if (q >= eol) break;
                            const char *qe = tokenEnd(q, eol);
                            verts[nTokens] = parseVertexFast(q, qe);
                            q = qe;
                        }
                        if (nTokens < 3) throw NoriException("Could not parse integer value \"\"");      /* parseVertex of an empty token */
                        int nVertices = 3;
                        if (nTokens == 4) { verts[4] = verts[0]; verts[5] = verts[2]; nVertices = 6; }
                        pc.corners.insert(pc.corners.end(), verts, verts + nVertices);
                    }
                    cur = eol + 1;
                }
            } catch (const std::exception &e) { pc.error = e.what(); }
        };
        {
            std::vector<std::thread> workers;
            for (size_t k = 1; k < nPieces; ++k) workers.emplace_back(readPiece, std::ref(pieces[k]));
            readPiece(pieces[0]);
            for (std::thread &w : workers) w.join();
        }
        lap("read + parse");
        for (Piece &pc : pieces) {                                      /* the first error in file order is the one a single pass reports */
            if (!pc.error.empty()) throw NoriException("%s", pc.error);
            positions.insert(positions.end(), pc.positions.begin(), pc.positions.end());
            normals.insert(normals.end(), pc.normals.begin(), pc.normals.end());
            texcoords.insert(texcoords.end(), pc.texcoords.begin(), pc.texcoords.end());
            m_bbox.expandBy(pc.bbox);
            corners.insert(corners.end(), pc.corners.begin(), pc.corners.end());
            pc = Piece();
        }
        /* corner numbering: nearly every corner of a mesh repeats the (uv, n) its position was first seen with — one array
         * indexed by position answers those; the rest go through the hash table */
        {
            std::vector<uint32_t> firstOf(positions.size() + 2, 0u);      /* vertex number + 1 of the first corner seen at a position */
            indices.reserve(corners.size());
            for (const OBJVertex &v : corners) {
                if ((size_t) v.p < firstOf.size()) {
                    uint32_t &f = firstOf[v.p];
                    if (f) {
                        const OBJVertex &o = vertices[f - 1];
                        if (o.uv == v.uv && o.n == v.n) { indices.push_back(f - 1); continue; }
                    } else {
                        /* first corner at this position: it is new unless the table already knows it — it cannot, a known corner
                         * would have set firstOf */
                        vertices.push_back(v);
                        f = (uint32_t) vertices.size();
                        indices.push_back(f - 1);
                        continue;
                    }
                }
                indices.push_back(lookup(v));
            }
        }
        lap("join + corner numbering");
        m_F = indices;
        const size_t nv = vertices.size();
        m_V.resize(nv * 3);
        for (size_t i = 0; i < nv; ++i) for (int k = 0; k < 3; ++k) m_V[3 * i + k] = positions.at(vertices[i].p - 1)[k];
        if (!normals.empty()) {
            m_N.resize(nv * 3);
            for (size_t i = 0; i < nv; ++i) for (int k = 0; k < 3; ++k) m_N[3 * i + k] = normals.at(vertices[i].n - 1)[k];
        }
        if (!texcoords.empty()) {
            m_UV.resize(nv * 2);
            for (size_t i = 0; i < nv; ++i) { m_UV[2 * i] = texcoords.at(vertices[i].uv - 1).x(); m_UV[2 * i + 1] = texcoords.at(vertices[i].uv - 1).y(); }
        }
        if (!normals.empty() && !texcoords.empty()) {                /* obj.cpp:183-229 */
            m_T.assign(nv * 3, 0.f); m_BT.assign(nv * 3, 0.f);
            auto add3 = [](std::vector<float> &a, uint32_t i, const Vector3f &v) { for (int k = 0; k < 3; ++k) a[3 * i + k] += v[k]; };
            for (size_t t = 0; t < indices.size() / 3; ++t) {
                const uint32_t i1 = indices[3 * t], i2 = indices[3 * t + 1], i3 = indices[3 * t + 2];
                const Vector3f v0 = vertex(i1), v1 = vertex(i2), v2 = vertex(i3);
                const Vector3f Edge1 = v1 - v0, Edge2 = v2 - v0;
                const float DeltaU1 = m_UV[2 * i2] - m_UV[2 * i1], DeltaV1 = m_UV[2 * i2 + 1] - m_UV[2 * i1 + 1];
                const float DeltaU2 = m_UV[2 * i3] - m_UV[2 * i1], DeltaV2 = m_UV[2 * i3 + 1] - m_UV[2 * i1 + 1];
                const float frac = DeltaU1 * DeltaV2 - DeltaU2 * DeltaV1;
                const float f = 1.f / (frac == 0.f ? 1e-4f : frac);
                const Vector3f Tangent(f * (DeltaV2 * Edge1.x() - DeltaV1 * Edge2.x()), f * (DeltaV2 * Edge1.y() - DeltaV1 * Edge2.y()),
                                       f * (DeltaV2 * Edge1.z() - DeltaV1 * Edge2.z()));
                const Vector3f tw = m_trafo.vector(Tangent);
                add3(m_T, i1, tw); add3(m_T, i2, tw); add3(m_T, i3, tw);
                const uint32_t ids[3] = { i1, i2, i3 };
                for (uint32_t id : ids) {
                    const Vector3f n(m_N[3 * id], m_N[3 * id + 1], m_N[3 * id + 2]);
                    add3(m_BT, id, m_trafo.vector(n.cross(Tangent)));
                }
            }
        }
        m_name = m_filename;
        lap("arrays");
    }
    std::string m_filename;
    Transform m_trafo;
};
NORI_REGISTER_CLASS(WavefrontOBJ, "obj")

} // namespace nori
