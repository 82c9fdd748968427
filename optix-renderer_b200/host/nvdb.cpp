/* nvdb.cpp — NanoVDB files (".nvdb") in and out, from the file layout: what the reference reads through
 * nanovdb::io::readGrid and writes as its cache through nanovdb::io::writeGrid (src/textures/NvdbVolume.vdb.cpp:9-81,
 * :120-133), with no NanoVDB library on this side. Layout of version 28.1.2 as the reference vendors it
 * (ext/nanovdb/nanovdb/NanoVDB.h, util/IO.h), float grids, uncompressed segments:
 *
 *   file   = segment*                                   segment = Header | (MetaData | name)* | grid blob*
 *   Header 16 B: magic "NanoVDB0", u16 major, minor, gridCount, codec (0 = none)
 *   MetaData 160 B: u64 gridSize, fileSize, nameKey, voxelCount; u32 gridType (1 = float), gridClass; f64 worldBBox[6];
 *            i32 indexBBox[6]; f64 voxelSize[3]; u32 nameSize; u32 nodeCount[4] (leaf, lower, upper, root); u16 codec, major
 *   blob   = GridData 672 B | TreeData 64 B | RootData 64 B | root tiles 32 B each | upper nodes | lower nodes | leaves
 *   GridData: magic @0, checksum @8, major @16, flags @20, gridSize @24, name[256] @32, Map @288 (matF, invMatF, vecF,
 *            taperF, matD @376, invMatD @448, vecD @520, taperD @544), worldBBox @552, voxelSize @600, class @624, type @628
 *   TreeData: u64 byte offsets from the tree to leaf / lower / upper / root arrays, u32 node counts[3]
 *   RootData: bbox i32[6], activeVoxelCount u64 @24, tileCount u32 @32, background, min, max, avg, stddev f32 @36..
 *   root tile: key u64 (z >> 12 | y >> 12 << 21 | x >> 12 << 42), childID i32 @8 (< 0: value tile), state u32, value f32 @16
 *   upper (32^3 tiles of 128^3) 139 328 B / lower (16^3 tiles of 8^3) 17 472 B: bbox i32[6], offset i32 @24, flags u32,
 *            value mask, child mask, min, max, avg, stddev, table of 4-byte entries, a union { f32 value | i32 childID }, at a 32-byte
 *            boundary; children are entries of the NEXT level's array
 *   leaf 2144 B: bboxMin i32[3], bboxDif u8[3], flags u8, value mask 64 B @16, min, max, avg, stddev @80, 512 values @96
 *   voxel index inside a node: x slowest, z fastest; mask bit n = word n >> 6, bit n & 63. */
#include "vdb.h"
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

namespace nori {

namespace {

const uint64_t NVDB_MAGIC = 0x304244566f6e614eULL;
enum { HEADER_BYTES = 16, META_BYTES = 160, GRID_BYTES = 672, TREE_BYTES = 64, ROOT_BYTES = 64, TILE_BYTES = 32,
       UPPER_BYTES = 139328, LOWER_BYTES = 17472, LEAF_BYTES = 2144,
       UPPER_TABLE = 8256, UPPER_CHILD_MASK = 4128, LOWER_TABLE = 1088, LOWER_CHILD_MASK = 544, LEAF_VALUES = 96, LEAF_MASK = 16 };

template <typename T> T rd(const uint8_t *p) { T v; memcpy(&v, p, sizeof(T)); return v; }
template <typename T> void wr(uint8_t *p, T v) { memcpy(p, &v, sizeof(T)); }
bool bit(const uint8_t *mask, uint32_t n) { return (rd<uint64_t>(mask + 8 * (n >> 6)) >> (n & 63)) & 1u; }
void setBit(uint8_t *mask, uint32_t n) { uint64_t w = rd<uint64_t>(mask + 8 * (n >> 6)); w |= 1ull << (n & 63); wr(mask + 8 * (n >> 6), w); }

struct Blob {                     /* one grid in memory */
    const uint8_t *grid = nullptr, *tree = nullptr, *root = nullptr, *upper = nullptr, *lower = nullptr, *leaf = nullptr;
    uint32_t nTiles = 0;
    std::map<uint64_t, const uint8_t *> tiles;
    float background = 0.f;

    /* ReadAccessor::getValue(Coord) of the reference, by walking the tree */
    float get(int x, int y, int z) const {
        const uint64_t key = (uint64_t) ((uint32_t) z >> 12) | ((uint64_t) ((uint32_t) y >> 12) << 21) | ((uint64_t) ((uint32_t) x >> 12) << 42);
        auto it = tiles.find(key);
        if (it == tiles.end()) return background;
        const uint8_t *t = it->second;
        const int32_t up = rd<int32_t>(t + 8);
        if (up < 0) return rd<float>(t + 16);
        const uint8_t *u = upper + (size_t) up * UPPER_BYTES;
        const uint32_t nu = (((uint32_t) x & 4095u) >> 7) << 10 | (((uint32_t) y & 4095u) >> 7) << 5 | (((uint32_t) z & 4095u) >> 7);
        if (!bit(u + UPPER_CHILD_MASK, nu)) return rd<float>(u + UPPER_TABLE + 4 * (size_t) nu);
        const uint8_t *l = lower + (size_t) rd<int32_t>(u + UPPER_TABLE + 4 * (size_t) nu) * LOWER_BYTES;
        const uint32_t nl = (((uint32_t) x & 127u) >> 3) << 8 | (((uint32_t) y & 127u) >> 3) << 4 | (((uint32_t) z & 127u) >> 3);
        if (!bit(l + LOWER_CHILD_MASK, nl)) return rd<float>(l + LOWER_TABLE + 4 * (size_t) nl);
        const uint8_t *f = leaf + (size_t) rd<int32_t>(l + LOWER_TABLE + 4 * (size_t) nl) * LEAF_BYTES;
        const uint32_t n = ((uint32_t) x & 7u) << 6 | ((uint32_t) y & 7u) << 3 | ((uint32_t) z & 7u);
        return rd<float>(f + LEAF_VALUES + 4 * (size_t) n);
    }
};

std::vector<uint8_t> readFile(const std::string &path) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) throw NoriException("cannot open %s", path);
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> data((size_t) std::max(0L, n));
    const bool ok = n >= 0 && fread(data.data(), 1, data.size(), f) == data.size();
    fclose(f);
    if (!ok) throw NoriException("cannot read %s", path);
    return data;
}

/* the grid's leaves as they are, plus lower-level value tiles (8^3 constant) that differ from the background as leaves of
 * their own; a 128^3 value tile that differs from the background has no place in a leaf list */
void readSparse(const Blob &b, uint32_t nUpper, uint32_t nLower, uint32_t nLeaf, VdbGrid &g, const std::string &path) {
    g.leaves.reserve(nLeaf);
    for (uint32_t n = 0; n < nLeaf; ++n) {
        const uint8_t *f = b.leaf + (size_t) n * LEAF_BYTES;
        VdbLeaf l;
        for (int a = 0; a < 3; ++a) l.org[a] = rd<int32_t>(f + 4 * a) & ~7;
        memcpy(l.v, f + LEAF_VALUES, sizeof(l.v));
        memcpy(l.mask, f + LEAF_MASK, sizeof(l.mask));
        g.leaves.push_back(l);
    }
    for (uint32_t n = 0; n < nLower; ++n) {
        const uint8_t *l = b.lower + (size_t) n * LOWER_BYTES;
        int org[3];
        for (int a = 0; a < 3; ++a) org[a] = rd<int32_t>(l + 4 * a) & ~127;
        for (uint32_t t = 0; t < 4096; ++t) {
            if (bit(l + LOWER_CHILD_MASK, t)) continue;
            const float v = rd<float>(l + LOWER_TABLE + 4 * (size_t) t);
            const bool active = bit(l + 32, t);
            if (v == b.background && !active) continue;
            VdbLeaf c;
            c.org[0] = org[0] + (int) (t >> 8) * 8; c.org[1] = org[1] + (int) ((t >> 4) & 15) * 8; c.org[2] = org[2] + (int) (t & 15) * 8;
            std::fill(c.v, c.v + 512, v);
            std::fill(c.mask, c.mask + 8, active ? ~0ull : 0ull);
            g.leaves.push_back(c);
        }
    }
    for (uint32_t n = 0; n < nUpper; ++n) {
        const uint8_t *u = b.upper + (size_t) n * UPPER_BYTES;
        for (uint32_t t = 0; t < 32768; ++t)
            if (!bit(u + UPPER_CHILD_MASK, t) && (rd<float>(u + UPPER_TABLE + 4 * (size_t) t) != b.background || bit(u + 32, t)))
                throw NoriException("%s: the grid holds a 128^3 value tile that differs from the background; not supported in the sparse form", path);
    }
    for (auto &t : b.tiles)
        if (rd<int32_t>(t.second + 8) < 0 && (rd<float>(t.second + 16) != b.background || rd<uint32_t>(t.second + 12)))
            throw NoriException("%s: the grid holds a 4096^3 value tile that differs from the background; not supported in the sparse form", path);
    g.indexLeaves();
}

} // namespace

VdbGrid vdbSparsify(const VdbGrid &src) {
    if (src.sparse()) return src;
    VdbGrid g = src;
    g.values.clear(); g.leaves.clear();
    g.active_voxels = 0; g.value_min = INFINITY; g.value_max = -INFINITY;
    int lo[3] = { INT32_MAX, INT32_MAX, INT32_MAX }, hi[3] = { INT32_MIN, INT32_MIN, INT32_MIN };
    if (!src.values.empty())
        for (int x = src.bbox_min[0] & ~7; x <= src.bbox_max[0]; x += 8)
            for (int y = src.bbox_min[1] & ~7; y <= src.bbox_max[1]; y += 8)
                for (int z = src.bbox_min[2] & ~7; z <= src.bbox_max[2]; z += 8) {
                    VdbLeaf l;
                    l.org[0] = x; l.org[1] = y; l.org[2] = z;
                    memset(l.mask, 0, sizeof(l.mask));
                    uint32_t active = 0;
                    for (int n = 0; n < 512; ++n) {
                        const int i = x + (n >> 6), j = y + ((n >> 3) & 7), k = z + (n & 7);
                        const float v = src.at(i, j, k);
                        l.v[n] = v;
                        if (v == src.background) continue;
                        l.mask[n >> 6] |= 1ull << (n & 63);
                        ++active;
                        g.value_min = std::min(g.value_min, v); g.value_max = std::max(g.value_max, v);
                        const int c[3] = { i, j, k };
                        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a]); hi[a] = std::max(hi[a], c[a]); }
                    }
                    if (active) { g.leaves.push_back(l); g.active_voxels += active; }
                }
    if (!g.active_voxels) { g.value_min = g.value_max = src.background; g.bbox_min[0] = g.bbox_min[1] = g.bbox_min[2] = 0; g.bbox_max[0] = g.bbox_max[1] = g.bbox_max[2] = -1; }
    else for (int a = 0; a < 3; ++a) { g.bbox_min[a] = lo[a]; g.bbox_max[a] = hi[a]; }
    for (int a = 0; a < 3; ++a) g.dims[a] = g.bbox_max[a] - g.bbox_min[a] + 1;
    g.indexLeaves();
    return g;
}

/* nanovdb::io::readGrid(file, gridName) (NvdbVolume.vdb.cpp:42-61): the named float grid — or the first one when the name is
 * empty — as the dense box the volume kernels sample: getValue over the index bounding box, background elsewhere. */
VdbGrid nvdbReadFloatGrid(const std::string &path, const std::string &gridName, bool keepSparse) {
    const std::vector<uint8_t> data = readFile(path);
    size_t pos = 0;
    while (pos + HEADER_BYTES <= data.size()) {
        const uint8_t *h = data.data() + pos;
        if (rd<uint64_t>(h) != NVDB_MAGIC) throw NoriException("%s: not a NanoVDB file (magic)", path);
        const uint16_t major = rd<uint16_t>(h + 8), count = rd<uint16_t>(h + 12), codec = rd<uint16_t>(h + 14);
        if (major != 28) throw NoriException("%s: NanoVDB file version %d, this reader knows the layout of version 28 (the reference's)", path, (int) major);
        pos += HEADER_BYTES;
        struct Meta { uint64_t gridSize, fileSize; uint32_t type, nameSize; std::string name; };
        std::vector<Meta> metas;
        for (uint16_t g = 0; g < count; ++g) {
            if (pos + META_BYTES > data.size()) throw NoriException("%s: truncated (grid meta data)", path);
            const uint8_t *m = data.data() + pos;
            Meta meta;
            meta.gridSize = rd<uint64_t>(m); meta.fileSize = rd<uint64_t>(m + 8); meta.type = rd<uint32_t>(m + 32); meta.nameSize = rd<uint32_t>(m + 136);
            pos += META_BYTES;
            if (pos + meta.nameSize > data.size()) throw NoriException("%s: truncated (grid name)", path);
            meta.name.assign((const char *) data.data() + pos, meta.nameSize ? meta.nameSize - 1 : 0);
            pos += meta.nameSize;
            metas.push_back(meta);
        }
        for (const Meta &meta : metas) {
            if (pos + meta.fileSize > data.size()) throw NoriException("%s: truncated (grid data)", path);
            const bool wanted = gridName.empty() || meta.name == gridName;
            if (wanted) {
                if (codec != 0) throw NoriException("%s: compressed NanoVDB segments (codec %d) are not supported; the reference writes its cache uncompressed", path, (int) codec);
                if (meta.type != 1) throw NoriException("%s: grid '%s' is not a float grid", path, meta.name);
                if (meta.gridSize < GRID_BYTES + TREE_BYTES + ROOT_BYTES) throw NoriException("%s: grid '%s' is too small to be a grid", path, meta.name);
                Blob b;
                b.grid = data.data() + pos;
                if (rd<uint64_t>(b.grid) != NVDB_MAGIC) throw NoriException("%s: grid '%s' has no NanoVDB magic number", path, meta.name);
                b.tree = b.grid + GRID_BYTES;
                const uint64_t oLeaf = rd<uint64_t>(b.tree), oLower = rd<uint64_t>(b.tree + 8), oUpper = rd<uint64_t>(b.tree + 16), oRoot = rd<uint64_t>(b.tree + 24);
                const uint32_t nLeaf = rd<uint32_t>(b.tree + 32), nLower = rd<uint32_t>(b.tree + 36), nUpper = rd<uint32_t>(b.tree + 40);
                const uint64_t room = meta.gridSize - GRID_BYTES;
                if (oRoot + ROOT_BYTES > room || oUpper + (uint64_t) nUpper * UPPER_BYTES > room || oLower + (uint64_t) nLower * LOWER_BYTES > room ||
                    oLeaf + (uint64_t) nLeaf * LEAF_BYTES > room)
                    throw NoriException("%s: grid '%s': node arrays outside the grid", path, meta.name);
                b.root = b.tree + oRoot; b.upper = b.tree + oUpper; b.lower = b.tree + oLower; b.leaf = b.tree + oLeaf;
                b.nTiles = rd<uint32_t>(b.root + 32);
                b.background = rd<float>(b.root + 36);
                if (oRoot + ROOT_BYTES + (uint64_t) b.nTiles * TILE_BYTES > room) throw NoriException("%s: grid '%s': root tiles outside the grid", path, meta.name);
                for (uint32_t t = 0; t < b.nTiles; ++t) {
                    const uint8_t *tile = b.root + ROOT_BYTES + (size_t) t * TILE_BYTES;
                    const int32_t child = rd<int32_t>(tile + 8);
                    if (child >= (int32_t) nUpper) throw NoriException("%s: grid '%s': root tile names a missing node", path, meta.name);
                    b.tiles[rd<uint64_t>(tile)] = tile;
                }
                VdbGrid g;
                g.name = meta.name; g.type = "NanoVDB float";
                for (int a = 0; a < 3; ++a) { g.bbox_min[a] = rd<int32_t>(b.root + 4 * a); g.bbox_max[a] = rd<int32_t>(b.root + 12 + 4 * a); }
                g.active_voxels = rd<uint64_t>(b.root + 24);
                g.background = b.background;
                g.value_min = rd<float>(b.root + 40); g.value_max = rd<float>(b.root + 44);
                const uint8_t *map = b.grid + 288;                       /* Map::applyMap: world = matD (row-major 3x3) * ijk + vecD */
                for (int r = 0; r < 3; ++r) {
                    for (int c = 0; c < 3; ++c) g.index_to_world[r * 4 + c] = rd<double>(map + 88 + 8 * (r * 3 + c));
                    g.index_to_world[r * 4 + 3] = rd<double>(map + 232 + 8 * r);
                    g.voxel_size[r] = rd<double>(b.grid + 600 + 8 * r);
                }
                bool empty = g.active_voxels == 0;
                for (int a = 0; a < 3; ++a) empty = empty || g.bbox_max[a] < g.bbox_min[a];
                if (empty) { g.bbox_max[0] = g.bbox_max[1] = g.bbox_max[2] = -1; g.bbox_min[0] = g.bbox_min[1] = g.bbox_min[2] = 0; return g; }
                const uint64_t vox = (uint64_t) (g.bbox_max[0] - g.bbox_min[0] + 1) * (uint64_t) (g.bbox_max[1] - g.bbox_min[1] + 1) * (uint64_t) (g.bbox_max[2] - g.bbox_min[2] + 1);
                for (int a = 0; a < 3; ++a) g.dims[a] = g.bbox_max[a] - g.bbox_min[a] + 1;
                /* sparse form: the file's leaves as they are, when the caller can take them and the box is mostly background
                 * or too large to fill (NB_VOLUME_SPARSE=1 / 0 forces the choice, for tests) */
                const char *force = getenv("NB_VOLUME_SPARSE");
                const bool wantSparse = force ? atoi(force) != 0 : (vox > (1ull << 27) || (uint64_t) nLeaf * 512 * 4 < vox);
                if (keepSparse && nLeaf > 0 && wantSparse) { readSparse(b, nUpper, nLower, nLeaf, g, path); return g; }
                if (vox > (1ull << 31)) throw NoriException("%s: grid '%s': a dense copy of its %llu-voxel box does not fit the volume kernels' layout", path, meta.name, (unsigned long long) vox);
                g.values.resize((size_t) vox);
                size_t n = 0;
                for (int i = 0; i < g.dims[0]; ++i)
                    for (int j = 0; j < g.dims[1]; ++j)
                        for (int k = 0; k < g.dims[2]; ++k) g.values[n++] = b.get(g.bbox_min[0] + i, g.bbox_min[1] + j, g.bbox_min[2] + k);
                return g;
            }
            pos += meta.fileSize;
        }
    }
    throw NoriException("%s: no float grid named '%s'", path, gridName);
}

namespace {

uint64_t nameHash(const std::string &s) {                 /* nanovdb::io::stringHash (util/IO.h:694-706) */
    uint64_t hash = 0;
    for (unsigned char c : s) { const uint64_t overflow = hash >> 56; hash *= 67; hash += c + overflow; }
    return hash;
}

struct Stats {                                            /* min / max / mean / population standard deviation of active values */
    float mn = INFINITY, mx = -INFINITY;
    double sum = 0.0, sum2 = 0.0;
    uint64_t n = 0;
    int lo[3] = { INT32_MAX, INT32_MAX, INT32_MAX }, hi[3] = { INT32_MIN, INT32_MIN, INT32_MIN };
    void add(float v, int x, int y, int z) {
        mn = std::min(mn, v); mx = std::max(mx, v); sum += v; sum2 += (double) v * v; ++n;
        const int c[3] = { x, y, z };
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a]); hi[a] = std::max(hi[a], c[a]); }
    }
    void merge(const Stats &o) {
        if (!o.n) return;
        mn = std::min(mn, o.mn); mx = std::max(mx, o.mx); sum += o.sum; sum2 += o.sum2; n += o.n;
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], o.lo[a]); hi[a] = std::max(hi[a], o.hi[a]); }
    }
    float avg() const { return n ? (float) (sum / (double) n) : 0.f; }
    float dev() const { if (!n) return 0.f; const double m = sum / (double) n; return (float) std::sqrt(std::max(0.0, sum2 / (double) n - m * m)); }
    void put(uint8_t *bboxAt, uint8_t *statsAt) const {    /* node header: bbox i32[6], then min, max, avg, stddev */
        for (int a = 0; a < 3; ++a) { wr<int32_t>(bboxAt + 4 * a, n ? lo[a] : 0); wr<int32_t>(bboxAt + 12 + 4 * a, n ? hi[a] : 0); }
        wr<float>(statsAt, n ? mn : 0.f); wr<float>(statsAt + 4, n ? mx : 0.f); wr<float>(statsAt + 8, avg()); wr<float>(statsAt + 12, dev());
    }
};

int floorTo(int v, int step) { return v & ~(step - 1); }     /* two's complement: also right for negative coordinates */

} // namespace

/* One float grid as a NanoVDB 28 file: the tree GridBuilder would build from the same voxels (every voxel of the dense box
 * active, everything else background), nodes in the order NanoVDB keeps them (upper, lower, leaf arrays, each sorted by
 * origin), uncompressed, no checksum. The reference's own reader (nanovdb::io::readGrid) reads it back voxel for voxel, with
 * the same bounding boxes, active count, min / max (tests/test_volume_cpu.py through oracle/_ref/nvdb_tool). */
void nvdbWriteFloatGrid(const VdbGrid &g, const std::string &path) {
    if (!g.sparse() && g.values.size() != (size_t) g.dims[0] * g.dims[1] * g.dims[2]) throw NoriException("nvdbWriteFloatGrid: grid '%s' is inconsistent", g.name);
    /* upper node (root tile key; tiles are kept sorted by key) -> lower node (its index in the upper) -> leaf (its index in the
     * lower): map order is array order at every level */
    struct Org { int v[3]; int32_t src; };                 /* src: index into g.leaves (sparse form) or -1 (fill from the dense box) */
    std::map<uint64_t, std::map<uint32_t, std::map<uint32_t, Org>>> tree;
    size_t nLeaves = 0;
    auto place = [&](int x, int y, int z, int32_t src) {
        const uint64_t key = (uint64_t) ((uint32_t) z >> 12) | ((uint64_t) ((uint32_t) y >> 12) << 21) | ((uint64_t) ((uint32_t) x >> 12) << 42);
        const uint32_t nu = (((uint32_t) x & 4095u) >> 7) << 10 | (((uint32_t) y & 4095u) >> 7) << 5 | (((uint32_t) z & 4095u) >> 7);
        const uint32_t nl = (((uint32_t) x & 127u) >> 3) << 8 | (((uint32_t) y & 127u) >> 3) << 4 | (((uint32_t) z & 127u) >> 3);
        tree[key][nu][nl] = { { x, y, z }, src };
        ++nLeaves;
    };
    if (g.sparse())
        for (size_t n = 0; n < g.leaves.size(); ++n) place(g.leaves[n].org[0], g.leaves[n].org[1], g.leaves[n].org[2], (int32_t) n);
    else if (!g.values.empty())
        for (int x = floorTo(g.bbox_min[0], 8); x <= g.bbox_max[0]; x += 8)
            for (int y = floorTo(g.bbox_min[1], 8); y <= g.bbox_max[1]; y += 8)
                for (int z = floorTo(g.bbox_min[2], 8); z <= g.bbox_max[2]; z += 8) place(x, y, z, -1);
    uint32_t nUpper = (uint32_t) tree.size(), nLower = 0, nLeaf = (uint32_t) nLeaves;
    for (auto &u : tree) nLower += (uint32_t) u.second.size();
    const uint32_t nTiles = nUpper;
    const uint64_t oRoot = TREE_BYTES, oUpper = oRoot + ROOT_BYTES + (uint64_t) nTiles * TILE_BYTES, oLower = oUpper + (uint64_t) nUpper * UPPER_BYTES,
                   oLeaf = oLower + (uint64_t) nLower * LOWER_BYTES, treeBytes = oLeaf + (uint64_t) nLeaf * LEAF_BYTES;
    const uint64_t gridSize = GRID_BYTES + treeBytes;
    std::vector<uint8_t> blob((size_t) gridSize, 0);
    uint8_t *grid = blob.data(), *tr = grid + GRID_BYTES, *root = tr + oRoot, *upper = tr + oUpper, *lower = tr + oLower, *leaf = tr + oLeaf;
    const int dy = g.dims[1], dz = g.dims[2];
    auto inside = [&](int x, int y, int z) {
        return x >= g.bbox_min[0] && x <= g.bbox_max[0] && y >= g.bbox_min[1] && y <= g.bbox_max[1] && z >= g.bbox_min[2] && z <= g.bbox_max[2];
    };
    Stats all;
    uint32_t iUpper = 0, iLower = 0, iLeaf = 0;
    for (auto &u : tree) {
        uint8_t *un = upper + (size_t) iUpper * UPPER_BYTES;
        Stats us;
        for (uint32_t n = 0; n < 32768; ++n) wr<float>(un + UPPER_TABLE + 4 * (size_t) n, g.background);
        for (auto &l : u.second) {
            uint8_t *ln = lower + (size_t) iLower * LOWER_BYTES;
            Stats ls;
            for (uint32_t n = 0; n < 4096; ++n) wr<float>(ln + LOWER_TABLE + 4 * (size_t) n, g.background);
            for (auto &f : l.second) {
                uint8_t *fn = leaf + (size_t) iLeaf * LEAF_BYTES;
                const int *o = f.second.v;
                Stats fs;
                const VdbLeaf *src = f.second.src >= 0 ? &g.leaves[(size_t) f.second.src] : nullptr;
                for (int n = 0; n < 512; ++n) {
                    const int x = o[0] + (n >> 6), y = o[1] + ((n >> 3) & 7), z = o[2] + (n & 7);
                    float v = g.background;
                    bool active = false;
                    if (src) { v = src->v[n]; active = (src->mask[n >> 6] >> (n & 63)) & 1u; }
                    else if (inside(x, y, z)) {
                        v = g.values[((size_t) (x - g.bbox_min[0]) * dy + (size_t) (y - g.bbox_min[1])) * dz + (size_t) (z - g.bbox_min[2])];
                        active = true;
                    }
                    if (active) { setBit(fn + LEAF_MASK, (uint32_t) n); fs.add(v, x, y, z); }
                    wr<float>(fn + LEAF_VALUES + 4 * (size_t) n, v);
                }
                if (!fs.n) for (int a = 0; a < 3; ++a) fs.lo[a] = fs.hi[a] = o[a];          /* a leaf without active voxels keeps its origin */
                for (int a = 0; a < 3; ++a) { wr<int32_t>(fn + 4 * a, fs.lo[a]); fn[12 + a] = (uint8_t) (fs.hi[a] - fs.lo[a]); }
                fn[15] = 2;                                                             /* LeafData::mFlags: the bounding box is set */
                wr<float>(fn + 80, fs.n ? fs.mn : g.background); wr<float>(fn + 84, fs.n ? fs.mx : g.background); wr<float>(fn + 88, fs.avg()); wr<float>(fn + 92, fs.dev());
                setBit(ln + LOWER_CHILD_MASK, f.first);
                wr<int32_t>(ln + LOWER_TABLE + 4 * (size_t) f.first, (int32_t) iLeaf);
                ls.merge(fs);
                ++iLeaf;
            }
            ls.put(ln, ln + LOWER_TABLE - 32);
            wr<int32_t>(ln + 24, (int32_t) (nLower - iLower));                         /* InternalData::mOffset: nodes from here to the child array */
            wr<uint32_t>(ln + 28, 2u);
            setBit(un + UPPER_CHILD_MASK, l.first);
            wr<int32_t>(un + UPPER_TABLE + 4 * (size_t) l.first, (int32_t) iLower);
            us.merge(ls);
            ++iLower;
        }
        us.put(un, un + UPPER_TABLE - 32);
        wr<int32_t>(un + 24, (int32_t) (nUpper - iUpper));
        wr<uint32_t>(un + 28, 2u);
        uint8_t *tile = root + ROOT_BYTES + (size_t) iUpper * TILE_BYTES;
        wr<uint64_t>(tile, u.first);
        wr<int32_t>(tile + 8, (int32_t) iUpper);
        all.merge(us);
        ++iUpper;
    }
    /* RootData, TreeData */
    for (int a = 0; a < 3; ++a) { wr<int32_t>(root + 4 * a, all.n ? all.lo[a] : INT32_MAX); wr<int32_t>(root + 12 + 4 * a, all.n ? all.hi[a] : INT32_MIN); }
    wr<uint64_t>(root + 24, all.n);
    wr<uint32_t>(root + 32, nTiles);
    wr<float>(root + 36, g.background);
    wr<float>(root + 40, all.n ? all.mn : g.background); wr<float>(root + 44, all.n ? all.mx : g.background);
    wr<float>(root + 48, all.avg()); wr<float>(root + 52, all.dev());
    wr<uint64_t>(tr, oLeaf); wr<uint64_t>(tr + 8, oLower); wr<uint64_t>(tr + 16, oUpper); wr<uint64_t>(tr + 24, oRoot);
    wr<uint32_t>(tr + 32, nLeaf); wr<uint32_t>(tr + 36, nLower); wr<uint32_t>(tr + 40, nUpper); wr<uint32_t>(tr + 44, 1u);
    wr<uint32_t>(tr + 48, nLower + nUpper + 1u); wr<uint32_t>(tr + 52, nUpper + 1u); wr<uint32_t>(tr + 56, 1u);   /* mPFSum: reversed prefix sum of the counts */
    /* GridData: Map = index_to_world and its inverse, in float and double (Map::set) */
    const double *M = g.index_to_world;
    const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], h = M[8], i = M[9], k = M[10];
    const double det = a * (e * k - f * i) - b * (d * k - f * h) + c * (d * i - e * h);
    if (!(std::fabs(det) > 0.0)) throw NoriException("nvdbWriteFloatGrid: grid '%s' has a singular index-to-world transform", g.name);
    const double fwd[9] = { a, b, c, d, e, f, h, i, k };
    const double inv[9] = { (e * k - f * i) / det, (c * i - b * k) / det, (b * f - c * e) / det, (f * h - d * k) / det, (a * k - c * h) / det,
                            (c * d - a * f) / det, (d * i - e * h) / det, (b * h - a * i) / det, (a * e - b * d) / det };
    auto apply = [&](double x, double y, double z, double *w) { for (int r = 0; r < 3; ++r) w[r] = M[r * 4] * x + M[r * 4 + 1] * y + M[r * 4 + 2] * z + M[r * 4 + 3]; };
    wr<uint64_t>(grid, NVDB_MAGIC);
    wr<uint64_t>(grid + 8, ~0ull);                                                      /* checksum: "not computed" (ChecksumMode::Disable) */
    wr<uint32_t>(grid + 16, 28u);
    wr<uint32_t>(grid + 20, all.n ? 30u : 28u);                                         /* HasBBox | HasMinMax | HasAverage | HasStdDeviation */
    wr<uint64_t>(grid + 24, gridSize);
    if (g.name.size() > 255) throw NoriException("nvdbWriteFloatGrid: grid name '%s' is longer than 255 characters", g.name);
    memcpy(grid + 32, g.name.data(), g.name.size());
    uint8_t *map = grid + 288;
    for (int n = 0; n < 9; ++n) {
        wr<float>(map + 4 * n, (float) fwd[n]); wr<float>(map + 36 + 4 * n, (float) inv[n]);
        wr<double>(map + 88 + 8 * n, fwd[n]); wr<double>(map + 160 + 8 * n, inv[n]);
    }
    for (int r = 0; r < 3; ++r) { wr<float>(map + 72 + 4 * r, (float) M[r * 4 + 3]); wr<double>(map + 232 + 8 * r, M[r * 4 + 3]); }
    wr<float>(map + 84, 1.f); wr<double>(map + 256, 1.0);
    double wlo[3] = { 0, 0, 0 }, whi[3] = { 0, 0, 0 };
    if (all.n)
        for (int corner = 0; corner < 8; ++corner) {                                    /* GridStats::processGrid: max is exclusive in world space */
            double w[3];
            apply(corner & 1 ? all.hi[0] + 1 : all.lo[0], corner & 2 ? all.hi[1] + 1 : all.lo[1], corner & 4 ? all.hi[2] + 1 : all.lo[2], w);
            for (int r = 0; r < 3; ++r) { wlo[r] = corner ? std::min(wlo[r], w[r]) : w[r]; whi[r] = corner ? std::max(whi[r], w[r]) : w[r]; }
        }
    double one[3], zero[3], voxel[3];
    apply(1, 1, 1, one); apply(0, 0, 0, zero);
    for (int r = 0; r < 3; ++r) { voxel[r] = one[r] - zero[r]; wr<double>(grid + 552 + 8 * r, wlo[r]); wr<double>(grid + 576 + 8 * r, whi[r]); wr<double>(grid + 600 + 8 * r, voxel[r]); }
    wr<uint32_t>(grid + 624, 2u);                                                       /* GridClass::FogVolume */
    wr<uint32_t>(grid + 628, 1u);                                                       /* GridType::Float */
    /* segment: Header | MetaData | name | grid */
    const uint32_t nameSize = (uint32_t) g.name.size() + 1;
    std::vector<uint8_t> head(HEADER_BYTES + META_BYTES + nameSize, 0);
    uint8_t *hd = head.data(), *m = hd + HEADER_BYTES;
    wr<uint64_t>(hd, NVDB_MAGIC);
    wr<uint16_t>(hd + 8, 28); wr<uint16_t>(hd + 10, 1); wr<uint16_t>(hd + 12, 1); wr<uint16_t>(hd + 14, 0);
    wr<uint64_t>(m, gridSize); wr<uint64_t>(m + 8, gridSize); wr<uint64_t>(m + 16, nameHash(g.name)); wr<uint64_t>(m + 24, all.n);
    wr<uint32_t>(m + 32, 1u); wr<uint32_t>(m + 36, 2u);
    for (int r = 0; r < 3; ++r) {
        wr<double>(m + 40 + 8 * r, wlo[r]); wr<double>(m + 64 + 8 * r, whi[r]);
        wr<int32_t>(m + 88 + 4 * r, rd<int32_t>(root + 4 * r)); wr<int32_t>(m + 100 + 4 * r, rd<int32_t>(root + 12 + 4 * r));
        wr<double>(m + 112 + 8 * r, voxel[r]);
    }
    wr<uint32_t>(m + 136, nameSize);
    wr<uint32_t>(m + 140, nLeaf); wr<uint32_t>(m + 144, nLower); wr<uint32_t>(m + 148, nUpper); wr<uint32_t>(m + 152, 1u);
    wr<uint16_t>(m + 156, 0); wr<uint16_t>(m + 158, 28);
    memcpy(m + META_BYTES, g.name.data(), g.name.size());
    FILE *out = fopen(path.c_str(), "wb");
    if (!out) throw NoriException("cannot write %s", path);
    const bool ok = fwrite(head.data(), 1, head.size(), out) == head.size() && fwrite(blob.data(), 1, blob.size(), out) == blob.size();
    if (fclose(out) != 0 || !ok) throw NoriException("cannot write %s (disk full?)", path);
}

} // namespace nori
