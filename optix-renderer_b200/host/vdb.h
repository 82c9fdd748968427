/* vdb.h — reader for OpenVDB .vdb files (what src/textures/NvdbVolume.vdb.cpp:83-134 does through
 * the OpenVDB library, absent here): the float grids of a file as dense voxel arrays over their
 * active bounding box, plus the index->world map. The reference converts such a grid to NanoVDB
 * and samples it by index coordinates; the dense array answers the same getValue(ijk) for every
 * ijk (background outside the box), which is all the volume kernels need. */
#pragma once
#include "nori.h"
#include <unordered_map>

namespace nori {

/* One 8^3 block of a sparse grid: origin (multiples of 8 in index space), values x slowest / z fastest, active-voxel mask
 * (bit n of word n >> 6 for voxel n) — a NanoVDB / OpenVDB leaf node. */
struct VdbLeaf {
    int org[3];
    float v[512];
    uint64_t mask[8];
};

struct VdbGrid {
    std::string name, type;                 /* e.g. "density", "Tree_float_5_4_3_HalfFloat" */
    int bbox_min[3] = { 0, 0, 0 }, bbox_max[3] = { -1, -1, -1 };     /* active voxels, inclusive (empty: max < min) */
    int dims[3] = { 0, 0, 0 };
    double index_to_world[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };   /* row-major 3x4 */
    double voxel_size[3] = { 1, 1, 1 };
    float background = 0.f;
    float value_max = 0.f, value_min = 0.f; /* over active voxels and active tiles (NanoVDB root().valueMax()) */
    uint64_t active_voxels = 0;
    /* DENSE form: every voxel of the box, dims[0]*dims[1]*dims[2], x slowest, z fastest (all of them count as active).
     * SPARSE form (values empty, leaves not): only the listed blocks, everything else background — what a .nvdb file holds,
     * kept as it is when the box is too large or too empty to fill (nvdbReadFloatGrid(.., keepSparse)). */
    std::vector<float> values;
    std::vector<VdbLeaf> leaves;
    std::unordered_map<uint64_t, uint32_t> leafIndex;
    bool sparse() const { return values.empty() && !leaves.empty(); }
    static uint64_t leafKey(int i, int j, int k) {
        /* 21 bits per axis: distinct for index coordinates within +-2^23 voxels, far beyond what the kernels' 32-bit array holds */
        return ((uint64_t) (((uint32_t) i >> 3) & 0x1fffffu) << 42) | ((uint64_t) (((uint32_t) j >> 3) & 0x1fffffu) << 21) | (uint64_t) (((uint32_t) k >> 3) & 0x1fffffu);
    }
    void indexLeaves() {
        leafIndex.clear();
        leafIndex.reserve(leaves.size() * 2);
        for (size_t n = 0; n < leaves.size(); ++n) leafIndex[leafKey(leaves[n].org[0], leaves[n].org[1], leaves[n].org[2])] = (uint32_t) n;
    }
    const VdbLeaf *leafAt(int i, int j, int k) const {
        auto it = leafIndex.find(leafKey(i, j, k));
        if (it == leafIndex.end()) return nullptr;
        const VdbLeaf &l = leaves[it->second];
        return (l.org[0] == (i & ~7) && l.org[1] == (j & ~7) && l.org[2] == (k & ~7)) ? &l : nullptr;
    }
    float at(int i, int j, int k) const {   /* ReadAccessor::getValue(Coord(i,j,k)) */
        if (sparse()) {
            const VdbLeaf *l = leafAt(i, j, k);
            return l ? l->v[(i & 7) << 6 | (j & 7) << 3 | (k & 7)] : background;
        }
        i -= bbox_min[0]; j -= bbox_min[1]; k -= bbox_min[2];
        if (i < 0 || j < 0 || k < 0 || i >= dims[0] || j >= dims[1] || k >= dims[2]) return background;
        return values[((size_t) i * dims[1] + j) * dims[2] + k];
    }
};

/* dense -> sparse: voxels that differ from the background become the active ones, blocks without any are dropped; bounding
 * box, active count, minimum and maximum follow (what a writer that only stores the smoke would have produced) */
VdbGrid vdbSparsify(const VdbGrid &g);

/* All grids' names and types, in file order (openvdb::io::File::beginName) */
std::vector<std::pair<std::string, std::string>> vdbListGrids(const std::string &path);
/* One float grid by name; throws NoriException if absent or not a float 5-4-3 tree */
VdbGrid vdbReadFloatGrid(const std::string &path, const std::string &gridName, bool keepSparse = false);   /* keepSparse: as nvdbReadFloatGrid */
/* ".nbv": a grid as a plain dense dump — "NBV1", i32 dims[3], i32 bbox_min[3], f64 index_to_world[12]
 * (row-major 3x4), f32 background, f32 values (x slowest). Every voxel of the box counts as active. */
void vdbWriteNbv(const VdbGrid &g, const std::string &path);
VdbGrid vdbReadNbv(const std::string &path);
/* ".nvdb": NanoVDB 28 files as the reference reads and caches them (nvdb.cpp). Reading takes the named float grid (the
 * first one when the name is empty); writing stores one float grid (dense form: every voxel of the box active; sparse form: the leaves and masks as they are), uncompressed —
 * the cache NvdbVolume::loadOpenVdbAndCacheNanoVdb leaves next to a .vdb (NvdbVolume.vdb.cpp:83-134). */
VdbGrid nvdbReadFloatGrid(const std::string &path, const std::string &gridName, bool keepSparse = false);
void nvdbWriteFloatGrid(const VdbGrid &g, const std::string &path);

}
