/* capi.cpp — a small C surface over the host library so the Python tests can drive it:
 * load a scene XML, flatten it (plugins' hooks + BVH build), write the result as .nbs. No GPU. */
#include <climits>
#include <cmath>
#include <cstring>
#include "plugins.h"
#include <memory>

static thread_local std::string g_error;

extern "C" {

const char *nbh_last_error(void) { return g_error.c_str(); }

/* XML -> flattened scene file. Returns 0, or -1 (error text in nbh_last_error), or -5 when a
 * plugin of the scene has no GPU kernel (NB_E_UNSUPPORTED). */
int nbh_flatten_xml_opts(const char *xml_path, const char *nbs_path, int host_tree);
int nbh_flatten_xml(const char *xml_path, const char *nbs_path) { return nbh_flatten_xml_opts(xml_path, nbs_path, 1); }
/* host_tree = 0: skip the host SAH build (seconds for a million triangles); the caller builds the tree on the device
 * (nb_build_bvh NB_BVH_DEVICE_SAH) and the file's bvh sections are empty */
int nbh_flatten_xml_opts(const char *xml_path, const char *nbs_path, int host_tree) {
    try {
        std::unique_ptr<nori::NoriObject> root(nori::loadFromXML(xml_path));
        if (root->getClassType() != nori::NoriObject::EScene) { g_error = "root element is not a <scene>"; return -1; }
        nori::FlatScene flat;
        std::string why;
        if (!static_cast<nori::Scene *>(root.get())->flatten(flat, &why, host_tree != 0)) { g_error = why; return NB_E_UNSUPPORTED; }
        flat.writeNbs(nbs_path);
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return -1;
    }
}

/* Parse only: reports the object tree's summary or the parser's error */
int nbh_parse_xml(const char *xml_path, char *summary, int summary_len) {
    try {
        std::unique_ptr<nori::NoriObject> root(nori::loadFromXML(xml_path));
        if (summary && summary_len > 0) snprintf(summary, summary_len, "%s", root->toString().c_str());
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return -1;
    }
}

}

/* ---- volume files (vdb.h) ------------------------------------------------------------------------ */
#include "vdb.h"
extern "C" {

/* "name:type\n" per grid */
int nbh_vdb_list(const char *path, char *out, int out_len) {
    try {
        std::string s;
        for (const auto &g : nori::vdbListGrids(path)) s += g.first + ":" + g.second + "\n";
        if (out && out_len > 0) snprintf(out, out_len, "%s", s.c_str());
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

/* One float grid as a dense array. bbox = {min xyz, max xyz} of the active voxels (inclusive),
 * stats = {background, min, max}. With values == NULL only the header is filled. Returns the
 * number of voxels of the dense array, or -1. */
long long nbh_vdb_read(const char *path, const char *grid, int32_t bbox[6], double index_to_world[12], float stats[3],
                       uint64_t *active_voxels, float *values, uint64_t capacity) {
    try {
        const std::string p(path);
        const bool nano = p.size() > 5 && p.substr(p.size() - 5) == ".nvdb";
        const nori::VdbGrid g = nano ? nori::nvdbReadFloatGrid(p, grid) : nori::vdbReadFloatGrid(p, grid);
        for (int k = 0; k < 3; ++k) { bbox[k] = g.bbox_min[k]; bbox[3 + k] = g.bbox_max[k]; }
        for (int k = 0; k < 12; ++k) index_to_world[k] = g.index_to_world[k];
        stats[0] = g.background; stats[1] = g.value_min; stats[2] = g.value_max;
        if (active_voxels) *active_voxels = g.active_voxels;
        if (values) {
            if (capacity < g.values.size()) { g_error = "nbh_vdb_read: buffer too small"; return -1; }
            std::copy(g.values.begin(), g.values.end(), values);
        }
        return (long long) g.values.size();
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

}

/* .vdb -> .nbv: one float grid as a dense dump ("NBV1", i32 dims[3], i32 bbox_min[3], f64
 * index_to_world[12], f32 background, f32 values x-slowest). The oracle harness reads it in place
 * of the OpenVDB library (oracle/ref_heterog.cpp); the `volume` plugin reads it too. */
extern "C" int nbh_vdb_to_nbv(const char *vdb_path, const char *grid, const char *nbv_path) {
    try {
        nori::vdbWriteNbv(nori::vdbReadFloatGrid(vdb_path, grid), nbv_path);
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

/* .vdb / .nbv / .nvdb -> .nvdb: the NanoVDB cache the reference would write (NvdbVolume.vdb.cpp:83-134) */
extern "C" int nbh_to_nvdb(const char *src_path, const char *grid, const char *nvdb_path, int sparse) {
    try {
        const std::string p(src_path);
        const std::string ext = p.rfind('.') == std::string::npos ? "" : p.substr(p.rfind('.'));
        /* sparse = 2: keep the source file's own leaves and masks (what openToNanoVDB keeps of a .vdb, NvdbVolume.vdb.cpp:118-119) */
        nori::VdbGrid g = ext == ".nbv" ? nori::vdbReadNbv(p) : ext == ".nvdb" ? nori::nvdbReadFloatGrid(p, grid, sparse == 2) : nori::vdbReadFloatGrid(p, grid, sparse == 2);
        if (g.name.empty() || ext == ".nbv") g.name = grid;          /* a .nbv dump carries no grid name */
        nori::nvdbWriteFloatGrid(sparse == 1 ? nori::vdbSparsify(g) : g, nvdb_path);    /* sparse: only blocks that hold something other than background */
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

/* A sparse float grid from a list of 8^3 blocks (origins multiples of 8, 512 values each, x slowest), written as .nvdb:
 * voxels that differ from the background are the active ones. index_to_world: row-major 3x4. */
extern "C" int nbh_nvdb_write_leaves(const char *nvdb_path, const char *grid, uint64_t n_leaves, const int32_t *origins, const float *values,
                                     const double *index_to_world, float background) {
    try {
        nori::VdbGrid g;
        g.name = grid; g.type = "NanoVDB float"; g.background = background;
        for (int i = 0; i < 12; ++i) g.index_to_world[i] = index_to_world[i];
        g.value_min = INFINITY; g.value_max = -INFINITY;
        int lo[3] = { INT32_MAX, INT32_MAX, INT32_MAX }, hi[3] = { INT32_MIN, INT32_MIN, INT32_MIN };
        for (uint64_t n = 0; n < n_leaves; ++n) {
            nori::VdbLeaf l;
            for (int a = 0; a < 3; ++a) {
                l.org[a] = origins[3 * n + a];
                if (l.org[a] & 7) throw nori::NoriException("nbh_nvdb_write_leaves: block %llu does not start at a multiple of 8", (unsigned long long) n);
            }
            memcpy(l.v, values + 512 * n, sizeof(l.v));
            memset(l.mask, 0, sizeof(l.mask));
            for (int v = 0; v < 512; ++v) {
                if (l.v[v] == background) continue;
                l.mask[v >> 6] |= 1ull << (v & 63);
                ++g.active_voxels;
                g.value_min = std::min(g.value_min, l.v[v]); g.value_max = std::max(g.value_max, l.v[v]);
                const int c[3] = { l.org[0] + (v >> 6), l.org[1] + ((v >> 3) & 7), l.org[2] + (v & 7) };
                for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a]); hi[a] = std::max(hi[a], c[a]); }
            }
            g.leaves.push_back(l);
        }
        if (!g.active_voxels) throw nori::NoriException("nbh_nvdb_write_leaves: no voxel differs from the background");
        for (int a = 0; a < 3; ++a) { g.bbox_min[a] = lo[a]; g.bbox_max[a] = hi[a]; g.dims[a] = hi[a] - lo[a] + 1; }
        g.indexLeaves();
        if (g.leafIndex.size() != g.leaves.size()) throw nori::NoriException("nbh_nvdb_write_leaves: two blocks share an origin");
        nori::nvdbWriteFloatGrid(g, nvdb_path);
        return 0;
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

/* ttest.cpp, for the tests: the tail probabilities the t-test / chi^2 harness decides with */
namespace nori { double studentTwoSidedP(double t, double dof); double chi2UpperP(double chsq, double dof); }
extern "C" double nbh_student_t_two_sided_p(double t, double dof) { return nori::studentTwoSidedP(t, dof); }
extern "C" double nbh_chi2_upper_p(double chsq, double dof) { return nori::chi2UpperP(chsq, dof); }

/* png.h, for the tests: the image writer of nori_b200 out.png */
#include "png.h"
extern "C" int nbh_write_png(const char *path, const float *rgb, uint32_t h, uint32_t w) {
    try { nori::writePng(path, rgb, h, w); return 0; } catch (const std::exception &e) { g_error = e.what(); return -1; }
}

/* codecs.h, for the tests: zlib stream -> bytes. Returns the decoded size or -1. */
#include "codecs.h"
extern "C" long long nbh_inflate(const uint8_t *src, uint64_t n, uint8_t *dst, uint64_t capacity) {
    try {
        const std::vector<uint8_t> out = nori::zlibInflate(src, (size_t) n);
        if (out.size() > capacity) { g_error = "nbh_inflate: buffer too small"; return -1; }
        std::copy(out.begin(), out.end(), dst);
        return (long long) out.size();
    } catch (const std::exception &e) { g_error = e.what(); return -1; }
}
