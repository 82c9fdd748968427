/* vdb.h — reader for OpenVDB .vdb files (what src/textures/NvdbVolume.vdb.cpp:83-134 does through
 * the OpenVDB library, absent here): the float grids of a file as dense voxel arrays over their
 * active bounding box, plus the index->world map. The reference converts such a grid to NanoVDB
 * and samples it by index coordinates; the dense array answers the same getValue(ijk) for every
 * ijk (background outside the box), which is all the volume kernels need. */
#pragma once
#include "nori.h"

namespace nori {

struct VdbGrid {
    std::string name, type;                 /* e.g. "density", "Tree_float_5_4_3_HalfFloat" */
    int bbox_min[3] = { 0, 0, 0 }, bbox_max[3] = { -1, -1, -1 };     /* active voxels, inclusive (empty: max < min) */
    int dims[3] = { 0, 0, 0 };
    double index_to_world[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };   /* row-major 3x4 */
    double voxel_size[3] = { 1, 1, 1 };
    float background = 0.f;
    float value_max = 0.f, value_min = 0.f; /* over active voxels and active tiles (NanoVDB root().valueMax()) */
    uint64_t active_voxels = 0;
    std::vector<float> values;              /* dims[0]*dims[1]*dims[2], x slowest, z fastest */
    float at(int i, int j, int k) const {   /* ReadAccessor::getValue(Coord(i,j,k)) */
        i -= bbox_min[0]; j -= bbox_min[1]; k -= bbox_min[2];
        if (i < 0 || j < 0 || k < 0 || i >= dims[0] || j >= dims[1] || k >= dims[2]) return background;
        return values[((size_t) i * dims[1] + j) * dims[2] + k];
    }
};

/* All grids' names and types, in file order (openvdb::io::File::beginName) */
std::vector<std::pair<std::string, std::string>> vdbListGrids(const std::string &path);
/* One float grid by name; throws NoriException if absent or not a float 5-4-3 tree */
VdbGrid vdbReadFloatGrid(const std::string &path, const std::string &gridName);
/* ".nbv": a grid as a plain dense dump — "NBV1", i32 dims[3], i32 bbox_min[3], f64 index_to_world[12]
 * (row-major 3x4), f32 background, f32 values (x slowest). Every voxel of the box counts as active. */
void vdbWriteNbv(const VdbGrid &g, const std::string &path);
VdbGrid vdbReadNbv(const std::string &path);
/* ".nvdb": NanoVDB 28 files as the reference reads and caches them (nvdb.cpp). Reading takes the named float grid (the
 * first one when the name is empty); writing stores one float grid, every voxel of the box active, uncompressed —
 * the cache NvdbVolume::loadOpenVdbAndCacheNanoVdb leaves next to a .vdb (NvdbVolume.vdb.cpp:83-134). */
VdbGrid nvdbReadFloatGrid(const std::string &path, const std::string &gridName);
void nvdbWriteFloatGrid(const VdbGrid &g, const std::string &path);

}
