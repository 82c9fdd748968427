/* vdb.cpp — OpenVDB file reader (float 5-4-3 trees, file format versions 222-224: per-grid
 * compression flags, node-mask compression, optional half-float storage, Blosc / zip / raw
 * buffers). Written from the published file layout; replaces the openvdb::io::File use of
 * src/textures/NvdbVolume.vdb.cpp:83-134. See vdb.h. */
#include <cstdlib>
#include "vdb.h"
#include "codecs.h"
#include <cstring>
#include <fstream>
#include <limits>

namespace nori {
namespace {

enum { COMPRESS_ZIP = 0x1, COMPRESS_ACTIVE_MASK = 0x2, COMPRESS_BLOSC = 0x4 };
enum { NO_MASK_OR_INACTIVE_VALS = 0, NO_MASK_AND_MINUS_BG, NO_MASK_AND_ONE_INACTIVE_VAL, MASK_AND_NO_INACTIVE_VALS,
       MASK_AND_ONE_INACTIVE_VAL, MASK_AND_TWO_INACTIVE_VALS, NO_MASK_AND_ALL_VALS };

struct Reader {
    std::vector<uint8_t> buf; size_t pos = 0; std::string path;
    const uint8_t *need(size_t n) {
        if (pos + n > buf.size() || pos + n < pos) throw NoriException("%s: unexpected end of file at byte %d", path, (int) pos);
        const uint8_t *p = buf.data() + pos; pos += n; return p;
    }
    template <typename T> T get() { T v; memcpy(&v, need(sizeof(T)), sizeof(T)); return v; }
    std::string str() { const uint32_t n = get<uint32_t>(); const uint8_t *p = need(n); return std::string((const char *) p, n); }
    void seek(size_t p) { if (p > buf.size()) throw NoriException("%s: offset %d beyond the file", path, (int) p); pos = p; }
};

float halfToFloat(uint16_t h) {
    const uint32_t s = (uint32_t) (h & 0x8000u) << 16, e = (h >> 10) & 31u, m = h & 1023u;
    uint32_t u;
    if (e == 0) {
        if (m == 0) u = s;
        else { int sh = 0; uint32_t mm = m; while (!(mm & 1024u)) { mm <<= 1; ++sh; } u = s | ((uint32_t) (113 - sh) << 23) | ((mm & 1023u) << 13); }
    } else if (e == 31) u = s | 0x7f800000u | (m << 13);
    else u = s | ((e + 112u) << 23) | (m << 13);
    float f; memcpy(&f, &u, 4); return f;
}

struct Mask {                               /* util::NodeMask: bit n = word n>>6, bit n&63 */
    std::vector<uint64_t> w; uint32_t bits = 0;
    void load(Reader &r, uint32_t nbits) { bits = nbits; w.resize(nbits / 64); memcpy(w.data(), r.need(nbits / 8), nbits / 8); }
    bool on(uint32_t n) const { return (w[n >> 6] >> (n & 63)) & 1u; }
    uint32_t countOn() const { uint32_t c = 0; for (uint64_t x : w) c += (uint32_t) __builtin_popcountll(x); return c; }
};

struct GridCtx { uint32_t compression = 0, version = 0; bool half = false; float background = 0.f; };

/* io::readData: `count` values of `size` bytes each */
std::vector<uint8_t> readData(Reader &r, const GridCtx &g, size_t count, size_t size) {
    const size_t bytes = count * size;
    std::vector<uint8_t> out;
    if (g.compression & (COMPRESS_BLOSC | COMPRESS_ZIP)) {
        const int64_t n = r.get<int64_t>();
        if (n <= 0) {                                        /* stored: -n raw bytes */
            const uint8_t *p = r.need((size_t) -n);
            out.assign(p, p + (size_t) -n);
        } else {
            const uint8_t *p = r.need((size_t) n);
            out = (g.compression & COMPRESS_BLOSC) ? bloscDecode(p, (size_t) n) : zlibInflate(p, (size_t) n, bytes);
        }
        if (out.size() < bytes) throw NoriException("%s: buffer decoded to %d bytes, %d expected", r.path, (int) out.size(), (int) bytes);
        out.resize(bytes);                                   /* Blosc input shorter than 128 bytes was zero-padded by the writer */
    } else {
        const uint8_t *p = r.need(bytes);
        out.assign(p, p + bytes);
    }
    return out;
}

/* io::readCompressedValues for float (optionally stored as half) */
void readCompressedValues(Reader &r, const GridCtx &g, float *dest, uint32_t destCount, const Mask &valueMask) {
    const bool maskCompressed = g.compression & COMPRESS_ACTIVE_MASK;
    int8_t metadata = NO_MASK_AND_ALL_VALS;
    if (g.version >= 222) metadata = r.get<int8_t>();
    float inactive1 = g.background, inactive0 = metadata == NO_MASK_OR_INACTIVE_VALS ? g.background : -g.background;
    if (metadata == NO_MASK_AND_ONE_INACTIVE_VAL || metadata == MASK_AND_ONE_INACTIVE_VAL || metadata == MASK_AND_TWO_INACTIVE_VALS) {
        inactive0 = r.get<float>();
        if (metadata == MASK_AND_TWO_INACTIVE_VALS) inactive1 = r.get<float>();
    }
    Mask selection;
    if (metadata == MASK_AND_NO_INACTIVE_VALS || metadata == MASK_AND_ONE_INACTIVE_VAL || metadata == MASK_AND_TWO_INACTIVE_VALS)
        selection.load(r, valueMask.bits);
    uint32_t count = destCount;
    if (maskCompressed && metadata != NO_MASK_AND_ALL_VALS && g.version >= 222) count = valueMask.countOn();
    std::vector<float> tmp(count);
    if (g.half) {                                            /* HalfReader::read returns before touching the stream when count < 1 */
        const std::vector<uint8_t> raw = count ? readData(r, g, count, 2) : std::vector<uint8_t>();
        for (uint32_t i = 0; i < count; ++i) { uint16_t h; memcpy(&h, raw.data() + 2 * i, 2); tmp[i] = halfToFloat(h); }
    } else {
        const std::vector<uint8_t> raw = readData(r, g, count, 4);
        if (count) memcpy(tmp.data(), raw.data(), (size_t) count * 4);
    }
    if (maskCompressed && count != destCount) {
        for (uint32_t i = 0, t = 0; i < destCount; ++i)
            dest[i] = valueMask.on(i) ? tmp[t++] : ((!selection.w.empty() && selection.on(i)) ? inactive1 : inactive0);
    } else
        for (uint32_t i = 0; i < destCount; ++i) dest[i] = tmp[i];
}

struct Leaf { int origin[3]; Mask valueMask; std::vector<float> values; };
struct Tile { int origin[3]; int dim; float value; bool active; };

struct Tree { std::vector<Leaf> leaves; std::vector<Tile> tiles; };

/* InternalNode<.., Log2Dim>::readTopology; childTotal = log2 of a child's extent */
void readInternal(Reader &r, const GridCtx &g, const int origin[3], int log2dim, int childTotal, Tree &tree) {
    const uint32_t n = 1u << (3 * log2dim);
    Mask childMask, valueMask;
    childMask.load(r, n);
    valueMask.load(r, n);
    std::vector<float> values(n);
    readCompressedValues(r, g, values.data(), n, valueMask);
    const uint32_t m = (1u << log2dim) - 1u;
    for (uint32_t i = 0; i < n; ++i) {
        const int o[3] = { origin[0] + (int) (((i >> (2 * log2dim)) & m) << childTotal), origin[1] + (int) (((i >> log2dim) & m) << childTotal),
                           origin[2] + (int) ((i & m) << childTotal) };
        if (childMask.on(i)) {
            if (childTotal == 3) {
                Leaf leaf; memcpy(leaf.origin, o, sizeof(o));
                leaf.valueMask.load(r, 512);                 /* LeafNode::readTopology */
                tree.leaves.push_back(std::move(leaf));
            } else
                readInternal(r, g, o, 4, 3, tree);
        } else if (valueMask.on(i) || values[i] != g.background)
            tree.tiles.push_back({ { o[0], o[1], o[2] }, 1 << childTotal, values[i], valueMask.on(i) });
    }
}

struct Descriptor { std::string name, type; int64_t gridPos, blockPos, endPos; };

struct FileHeader { uint32_t version; bool hasOffsets; std::vector<Descriptor> grids; };

void skipMetadata(Reader &r) {
    const int32_t n = r.get<int32_t>();
    for (int32_t i = 0; i < n; ++i) { r.str(); r.str(); const uint32_t size = r.get<uint32_t>(); r.need(size); }
}

Reader openFile(const std::string &path) {
    Reader r; r.path = path;
    std::ifstream f(path, std::ios::binary);
    if (!f) throw NoriException("Unable to open VDB file \"%s\"!", path);
    r.buf.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    return r;
}

FileHeader readHeader(Reader &r) {
    FileHeader h;
    if (r.get<int64_t>() != 0x56444220) throw NoriException("%s: not a VDB file", r.path);
    h.version = r.get<uint32_t>();
    if (h.version < 222) throw NoriException("%s: VDB file format %d is older than 222 and not supported", r.path, (int) h.version);
    r.get<uint32_t>(); r.get<uint32_t>();                    /* library version */
    h.hasOffsets = r.get<uint8_t>() != 0;
    r.need(36);                                              /* UUID */
    skipMetadata(r);
    const int32_t n = r.get<int32_t>();
    if (!h.hasOffsets && n > 1) throw NoriException("%s: streamed VDB files with several grids are not supported", r.path);
    for (int32_t i = 0; i < n; ++i) {
        Descriptor d;
        d.name = r.str(); d.type = r.str();
        r.str();                                             /* instance parent */
        d.gridPos = r.get<int64_t>(); d.blockPos = r.get<int64_t>(); d.endPos = r.get<int64_t>();
        const size_t sep = d.name.find('\x1e');              /* unique-name suffix */
        if (sep != std::string::npos) d.name.resize(sep);
        h.grids.push_back(d);
        if (h.hasOffsets) r.seek((size_t) d.endPos); else break;
    }
    return h;
}

void readTransform(Reader &r, VdbGrid &out) {
    const std::string type = r.str();
    double *M = out.index_to_world;
    auto vec3 = [&](double v[3]) { for (int i = 0; i < 3; ++i) v[i] = r.get<double>(); };
    double t[3] = { 0, 0, 0 }, s[3] = { 1, 1, 1 }, skip[3];
    if (type == "UniformScaleMap" || type == "ScaleMap") { vec3(s); for (int i = 0; i < 4; ++i) vec3(skip); }
    else if (type == "UniformScaleTranslateMap" || type == "ScaleTranslateMap") { vec3(t); vec3(s); for (int i = 0; i < 4; ++i) vec3(skip); }
    else if (type == "TranslationMap") vec3(t);
    else if (type == "AffineMap" || type == "UnitaryMap") {  /* row-vector convention: p' = p * A */
        double A[16];
        for (int i = 0; i < 16; ++i) A[i] = r.get<double>();
        for (int row = 0; row < 3; ++row) {
            for (int c = 0; c < 3; ++c) M[row * 4 + c] = A[c * 4 + row];
            M[row * 4 + 3] = A[12 + row];
            out.voxel_size[row] = std::sqrt(A[row * 4] * A[row * 4] + A[row * 4 + 1] * A[row * 4 + 1] + A[row * 4 + 2] * A[row * 4 + 2]);
        }
        return;
    } else
        throw NoriException("%s: transform map \"%s\" is not supported", r.path, type);
    for (int i = 0; i < 12; ++i) M[i] = 0.0;
    for (int i = 0; i < 3; ++i) { M[i * 4 + i] = s[i]; M[i * 4 + 3] = t[i]; out.voxel_size[i] = std::fabs(s[i]); }
}

}  // namespace

std::vector<std::pair<std::string, std::string>> vdbListGrids(const std::string &path) {
    Reader r = openFile(path);
    std::vector<std::pair<std::string, std::string>> out;
    for (const Descriptor &d : readHeader(r).grids) out.push_back({ d.name, d.type });
    return out;
}

VdbGrid vdbReadFloatGrid(const std::string &path, const std::string &gridName, bool keepSparse) {
    Reader r = openFile(path);
    const FileHeader h = readHeader(r);
    const Descriptor *d = nullptr;
    for (const Descriptor &c : h.grids) if (c.name == gridName) { d = &c; break; }
    if (!d) throw NoriException("%s: no grid named \"%s\"", path, gridName);
    GridCtx g; g.version = h.version;
    if (d->type == "Tree_float_5_4_3_HalfFloat") g.half = true;
    else if (d->type != "Tree_float_5_4_3") throw NoriException("%s: grid \"%s\" has type %s, only float 5-4-3 trees are supported", path, gridName, d->type);
    VdbGrid out; out.name = d->name; out.type = d->type;
    if (h.hasOffsets) r.seek((size_t) d->gridPos);
    g.compression = r.get<uint32_t>();
    skipMetadata(r);
    readTransform(r, out);
    /* Tree::readTopology */
    if (r.get<int32_t>() != 1) throw NoriException("%s: multi-buffer trees are not supported", path);
    g.background = out.background = r.get<float>();
    const uint32_t numTiles = r.get<uint32_t>(), numChildren = r.get<uint32_t>();
    Tree tree;
    for (uint32_t i = 0; i < numTiles; ++i) {
        Tile t; for (int k = 0; k < 3; ++k) t.origin[k] = r.get<int32_t>();
        t.value = r.get<float>(); t.active = r.get<uint8_t>() != 0; t.dim = 4096;
        if (t.active || t.value != g.background) tree.tiles.push_back(t);
    }
    for (uint32_t i = 0; i < numChildren; ++i) {
        int o[3]; for (int k = 0; k < 3; ++k) o[k] = r.get<int32_t>();
        readInternal(r, g, o, 5, 7, tree);
    }
    /* Tree::readBuffers: leaves in the same order */
    if (h.hasOffsets) r.seek((size_t) d->blockPos);
    for (Leaf &leaf : tree.leaves) {
        leaf.valueMask.load(r, 512);
        leaf.values.resize(512);
        readCompressedValues(r, g, leaf.values.data(), 512, leaf.valueMask);
    }
    /* active bounding box, extrema, dense array over everything that holds a non-background value */
    const int big = std::numeric_limits<int>::max();
    int amin[3] = { big, big, big }, amax[3] = { -big, -big, -big }, smin[3] = { big, big, big }, smax[3] = { -big, -big, -big };
    float vmax = -std::numeric_limits<float>::infinity(), vmin = std::numeric_limits<float>::infinity();
    auto grow = [](int lo[3], int hi[3], const int p[3]) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } };
    for (const Leaf &leaf : tree.leaves) {
        const int hi[3] = { leaf.origin[0] + 7, leaf.origin[1] + 7, leaf.origin[2] + 7 };
        grow(smin, smax, leaf.origin); grow(smin, smax, hi);
        for (uint32_t n = 0; n < 512; ++n) if (leaf.valueMask.on(n)) {
            const int p[3] = { leaf.origin[0] + (int) (n >> 6), leaf.origin[1] + (int) ((n >> 3) & 7), leaf.origin[2] + (int) (n & 7) };
            grow(amin, amax, p);
            vmax = std::max(vmax, leaf.values[n]); vmin = std::min(vmin, leaf.values[n]);
            ++out.active_voxels;
        }
    }
    for (const Tile &t : tree.tiles) {
        const int hi[3] = { t.origin[0] + t.dim - 1, t.origin[1] + t.dim - 1, t.origin[2] + t.dim - 1 };
        grow(smin, smax, t.origin); grow(smin, smax, hi);
        if (t.active) {
            grow(amin, amax, t.origin); grow(amin, amax, hi);
            vmax = std::max(vmax, t.value); vmin = std::min(vmin, t.value);
            out.active_voxels += (uint64_t) t.dim * t.dim * t.dim;
        }
    }
    if (out.active_voxels == 0) { out.value_max = out.value_min = out.background; return out; }
    out.value_max = vmax; out.value_min = vmin;
    /* the dense array covers the active box; stored values outside it (inactive voxels of
     * boundary leaves) equal the background in fog volumes and are dropped */
    double total = 1.0;
    for (int k = 0; k < 3; ++k) { out.bbox_min[k] = amin[k]; out.bbox_max[k] = amax[k]; out.dims[k] = amax[k] - amin[k] + 1; total *= out.dims[k]; }
    /* sparse form (vdb.h): the file's leaves with their masks — what openToNanoVDB keeps (NvdbVolume.vdb.cpp:118-119) — when the
     * caller can take them and the box is mostly background or too large to fill; 8^3 tiles become constant leaves */
    {
        const char *force = getenv("NB_VOLUME_SPARSE");
        const bool wantSparse = force ? atoi(force) != 0 : (total > 134217728.0 || (double) tree.leaves.size() * 512.0 * 4.0 < total);
        if (keepSparse && wantSparse && !tree.leaves.empty()) {
            for (const Tile &t : tree.tiles)
                if (t.dim > 8) throw NoriException("%s: grid \"%s\" holds a %d^3 value tile; not supported in the sparse form", path, gridName, t.dim);
            out.leaves.reserve(tree.leaves.size() + tree.tiles.size());
            for (const Leaf &leaf : tree.leaves) {
                VdbLeaf l;
                memcpy(l.org, leaf.origin, sizeof(l.org));
                memcpy(l.v, leaf.values.data(), sizeof(l.v));
                for (int w = 0; w < 8; ++w) l.mask[w] = 0;
                for (uint32_t n = 0; n < 512; ++n) if (leaf.valueMask.on(n)) l.mask[n >> 6] |= 1ull << (n & 63);
                out.leaves.push_back(l);
            }
            for (const Tile &t : tree.tiles) {
                VdbLeaf l;
                memcpy(l.org, t.origin, sizeof(l.org));
                std::fill(l.v, l.v + 512, t.value);
                std::fill(l.mask, l.mask + 8, t.active ? ~0ull : 0ull);
                out.leaves.push_back(l);
            }
            out.indexLeaves();
            return out;
        }
    }
    if (total > 2147483648.0) throw NoriException("%s: grid \"%s\" spans %dx%dx%d voxels, too large for a dense array", path, gridName, out.dims[0], out.dims[1], out.dims[2]);
    out.values.assign((size_t) total, out.background);
    auto put = [&](int i, int j, int k, float v) {
        i -= amin[0]; j -= amin[1]; k -= amin[2];
        if (i < 0 || j < 0 || k < 0 || i >= out.dims[0] || j >= out.dims[1] || k >= out.dims[2]) return;
        out.values[((size_t) i * out.dims[1] + j) * out.dims[2] + k] = v;
    };
    for (const Tile &t : tree.tiles)
        for (int i = std::max(t.origin[0], amin[0]); i <= std::min(t.origin[0] + t.dim - 1, amax[0]); ++i)
            for (int j = std::max(t.origin[1], amin[1]); j <= std::min(t.origin[1] + t.dim - 1, amax[1]); ++j)
                for (int k = std::max(t.origin[2], amin[2]); k <= std::min(t.origin[2] + t.dim - 1, amax[2]); ++k) put(i, j, k, t.value);
    for (const Leaf &leaf : tree.leaves)
        for (uint32_t n = 0; n < 512; ++n)
            put(leaf.origin[0] + (int) (n >> 6), leaf.origin[1] + (int) ((n >> 3) & 7), leaf.origin[2] + (int) (n & 7), leaf.values[n]);
    return out;
}

void vdbWriteNbv(const VdbGrid &g, const std::string &path) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw NoriException("cannot write %s", path);
    const int32_t dims[3] = { g.dims[0], g.dims[1], g.dims[2] }, org[3] = { g.bbox_min[0], g.bbox_min[1], g.bbox_min[2] };
    fwrite("NBV1", 1, 4, f); fwrite(dims, 4, 3, f); fwrite(org, 4, 3, f); fwrite(g.index_to_world, 8, 12, f); fwrite(&g.background, 4, 1, f);
    fwrite(g.values.data(), 4, g.values.size(), f);
    fclose(f);
}

VdbGrid vdbReadNbv(const std::string &path) {
    Reader r = openFile(path);
    if (memcmp(r.need(4), "NBV1", 4) != 0) throw NoriException("%s: not a .nbv volume", path);
    VdbGrid g; g.name = "density"; g.type = "nbv";
    double total = 1.0;
    for (int k = 0; k < 3; ++k) { g.dims[k] = r.get<int32_t>(); if (g.dims[k] <= 0) throw NoriException("%s: empty grid", path); total *= g.dims[k]; }
    for (int k = 0; k < 3; ++k) { g.bbox_min[k] = r.get<int32_t>(); g.bbox_max[k] = g.bbox_min[k] + g.dims[k] - 1; }
    for (int k = 0; k < 12; ++k) g.index_to_world[k] = r.get<double>();
    for (int k = 0; k < 3; ++k) g.voxel_size[k] = std::sqrt(g.index_to_world[k] * g.index_to_world[k] + g.index_to_world[4 + k] * g.index_to_world[4 + k] + g.index_to_world[8 + k] * g.index_to_world[8 + k]);
    g.background = r.get<float>();
    if (total > 2147483648.0) throw NoriException("%s: grid too large", path);
    g.values.resize((size_t) total);
    memcpy(g.values.data(), r.need(g.values.size() * 4), g.values.size() * 4);
    g.active_voxels = g.values.size();
    g.value_max = g.value_min = g.values[0];
    for (float v : g.values) { g.value_max = std::max(g.value_max, v); g.value_min = std::min(g.value_min, v); }
    return g;
}

}
