/* mesh.cpp — Mesh (src/shapes/mesh.cpp) and the Wavefront OBJ loader (src/shapes/obj.cpp:77-238):
 * toWorld baked into positions/normals while reading, (v, vt, vn) triples deduplicated in order
 * of first use, quads split 0-1-2 / 3-0-2, per-vertex tangents/bitangents accumulated per
 * triangle; area CDF as DiscretePDF builds it (include/nori/dpdf.h:40-113). */
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
    void load() {
        std::ifstream is(m_filename);
        if (is.fail()) throw NoriException("Unable to open OBJ file \"%s\"!", m_filename);
        std::vector<Vector3f> positions, normals;
        std::vector<Vector2f> texcoords;
        std::vector<uint32_t> indices;
        std::vector<OBJVertex> vertices;
        std::map<std::tuple<uint32_t, uint32_t, uint32_t>, uint32_t> vertexMap;
        std::string line_str;
        while (std::getline(is, line_str)) {
            std::istringstream line(line_str);
            std::string prefix;
            line >> prefix;
            if (prefix == "v") {
                float x = 0, y = 0, z = 0;
                line >> x >> y >> z;
                Point3f p = m_trafo.point(Point3f(x, y, z));
                m_bbox.expandBy(p);
                positions.push_back(p);
            } else if (prefix == "vt") {
                float x = 0, y = 0;
                line >> x >> y;
                texcoords.push_back(Vector2f(x, y));
            } else if (prefix == "vn") {
                float x = 0, y = 0, z = 0;
                line >> x >> y >> z;
                normals.push_back(m_trafo.normal(Normal3f(x, y, z)).normalized());
            } else if (prefix == "f") {
                std::string v1, v2, v3, v4;
                line >> v1 >> v2 >> v3 >> v4;
                OBJVertex verts[6];
                int nVertices = 3;
                verts[0] = parseVertex(v1); verts[1] = parseVertex(v2); verts[2] = parseVertex(v3);
                if (!v4.empty()) { verts[3] = parseVertex(v4); verts[4] = verts[0]; verts[5] = verts[2]; nVertices = 6; }
                for (int i = 0; i < nVertices; ++i) {
                    const OBJVertex &v = verts[i];
                    auto key = std::make_tuple(v.p, v.uv, v.n);
                    auto it = vertexMap.find(key);
                    if (it == vertexMap.end()) {
                        vertexMap[key] = (uint32_t) vertices.size();
                        indices.push_back((uint32_t) vertices.size());
                        vertices.push_back(v);
                    } else indices.push_back(it->second);
                }
            }
        }
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
    }
    std::string m_filename;
    Transform m_trafo;
};
NORI_REGISTER_CLASS(WavefrontOBJ, "obj")

} // namespace nori
