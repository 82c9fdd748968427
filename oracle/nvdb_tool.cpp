/* oracle/nvdb_tool.cpp — TEST INFRASTRUCTURE, not product.
 *
 * The reference's own NanoVDB (ext/nanovdb 28.1.2, header only) as the checker of host/nvdb.cpp:
 *   nvdb_tool write in.nbv out.nvdb [name]   builds a float grid from a dense .nbv dump with nanovdb::GridBuilder and
 *                                            writes it with nanovdb::io::writeGrid — the file the reference's
 *                                            NvdbVolume.vdb.cpp:15-39 caches and reads (fixtures for the reader)
 *   nvdb_tool dump in.nvdb out.bin [grid]    nanovdb::io::readGrid + ReadAccessor::getValue over the index bounding box
 *                                            widened by 2 voxels: i32 min[3], i32 dims[3], f64 index_to_world[12],
 *                                            f32 background, min, max, u64 active voxels, then the values (x slowest)
 *                                            — what the reference would see in a file written by host/nvdb.cpp
 *   nvdb_tool info in.nvdb                   the same reader on a grid too large to dump: box, active count, getValue probes
 */
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <nanovdb/NanoVDB.h>
#include <nanovdb/util/GridBuilder.h>
#include <nanovdb/util/IO.h>

static int fail(const char *what) { fprintf(stderr, "nvdb_tool: %s\n", what); return 1; }

static int modeWrite(int argc, char **argv) {
    if (argc < 4) return 2;
    FILE *f = fopen(argv[2], "rb");
    if (!f) return fail("cannot open the .nbv file");
    char magic[4]; int32_t dims[3], org[3]; double M[12]; float background;
    bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "NBV1", 4) == 0 && fread(dims, 4, 3, f) == 3 && fread(org, 4, 3, f) == 3 &&
              fread(M, 8, 12, f) == 12 && fread(&background, 4, 1, f) == 1;
    std::vector<float> values;
    if (ok) { values.resize((size_t) dims[0] * dims[1] * dims[2]); ok = fread(values.data(), 4, values.size(), f) == values.size(); }
    fclose(f);
    if (!ok) return fail("not a complete .nbv file");
    nanovdb::GridBuilder<float> builder(background);
    {
        auto acc = builder.getAccessor();
        size_t n = 0;
        for (int i = 0; i < dims[0]; ++i)
            for (int j = 0; j < dims[1]; ++j)
                for (int k = 0; k < dims[2]; ++k) acc.setValue(nanovdb::Coord(org[0] + i, org[1] + j, org[2] + k), values[n++]);
    }
    const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], g = M[6], h = M[8], i = M[9], k = M[10];
    const double det = a * (e * k - g * i) - b * (d * k - g * h) + c * (d * i - e * h);
    const double inv[3][3] = { { (e * k - g * i) / det, (c * i - b * k) / det, (b * g - c * e) / det },
                               { (g * h - d * k) / det, (a * k - c * h) / det, (c * d - a * g) / det },
                               { (d * i - e * h) / det, (b * h - a * i) / det, (a * e - b * d) / det } };
    double mat[4][4] = {}, invMat[4][4] = {};
    for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 3; ++cc) { mat[cc][r] = M[r * 4 + cc]; invMat[cc][r] = inv[r][cc]; }
    for (int r = 0; r < 3; ++r) { mat[3][r] = M[r * 4 + 3]; invMat[3][r] = -M[r * 4 + 3]; }
    mat[3][3] = invMat[3][3] = 1.0;
    nanovdb::Map map;
    map.set(mat, invMat, 1.0);
    auto handle = builder.getHandle<nanovdb::HostBuffer>(map, argc > 4 ? argv[4] : "density", nanovdb::GridClass::FogVolume);
    nanovdb::io::writeGrid(argv[3], handle);
    const auto *grid = handle.grid<float>();
    printf("{\"mode\": \"write\", \"bytes\": %llu, \"active\": %llu, \"max\": %.9g}\n", (unsigned long long) handle.size(),
           (unsigned long long) grid->activeVoxelCount(), (double) grid->tree().root().valueMax());
    return 0;
}

static int modeDump(int argc, char **argv) {
    if (argc < 4) return 2;
    nanovdb::GridHandle<nanovdb::HostBuffer> handle;
    if (argc > 4) handle = nanovdb::io::readGrid(argv[2], std::string(argv[4]));
    else handle = nanovdb::io::readGrid(argv[2], 0);
    const auto *grid = handle.grid<float>();
    if (!grid) return fail("no float grid in the file");
    const nanovdb::CoordBBox box = grid->indexBBox();
    const int B = 2;
    int32_t mn[3], dims[3];
    for (int a = 0; a < 3; ++a) { mn[a] = box.min()[a] - B; dims[a] = box.max()[a] - box.min()[a] + 1 + 2 * B; }
    std::vector<float> values((size_t) dims[0] * dims[1] * dims[2]);
    auto acc = grid->getAccessor();
    size_t n = 0;
    for (int i = 0; i < dims[0]; ++i)
        for (int j = 0; j < dims[1]; ++j)
            for (int k = 0; k < dims[2]; ++k) values[n++] = acc.getValue(nanovdb::Coord(mn[0] + i, mn[1] + j, mn[2] + k));
    double M[12];
    const nanovdb::Vec3d o = grid->indexToWorld(nanovdb::Vec3d(0.0));
    for (int c = 0; c < 3; ++c) {
        nanovdb::Vec3d e(0.0); e[c] = 1.0;
        const nanovdb::Vec3d w = grid->indexToWorld(e);
        for (int r = 0; r < 3; ++r) M[r * 4 + c] = w[r] - o[r];
    }
    for (int r = 0; r < 3; ++r) M[r * 4 + 3] = o[r];
    const float stats[3] = { grid->tree().background(), grid->tree().root().valueMin(), grid->tree().root().valueMax() };
    const uint64_t active = grid->activeVoxelCount();
    FILE *f = fopen(argv[3], "wb");
    if (!f) return fail("cannot write the dump");
    fwrite(mn, 4, 3, f); fwrite(dims, 4, 3, f); fwrite(M, 8, 12, f); fwrite(stats, 4, 3, f); fwrite(&active, 8, 1, f);
    fwrite(values.data(), 4, values.size(), f);
    fclose(f);
    printf("{\"mode\": \"dump\", \"name\": \"%s\", \"active\": %llu, \"min\": [%d, %d, %d], \"dims\": [%d, %d, %d], \"max\": %.9g}\n", grid->gridName(),
           (unsigned long long) active, mn[0], mn[1], mn[2], dims[0], dims[1], dims[2], (double) stats[2]);
    return 0;
}

/* info in.nvdb: what nanovdb::io::readGrid sees of a grid too large to dump — bounding box, active count, and getValue at
 * voxel (1, 2, 3) of the leaves that hold the box's minimum corner, its middle and its maximum corner, and at one empty place */
static int modeInfo(int argc, char **argv) {
    if (argc < 3) return 2;
    auto handle = nanovdb::io::readGrid(argv[2], 0);
    const auto *grid = handle.grid<float>();
    if (!grid) return fail("no float grid in the file");
    const nanovdb::CoordBBox box = grid->indexBBox();
    auto acc = grid->getAccessor();
    float probe[4];
    int n = 0;
    for (int leaf = 0; leaf < (int) grid->tree().nodeCount(0) && n < 3; ++leaf) {
        const auto *l = grid->tree().getNode<0>(leaf);
        const nanovdb::Coord o = l->origin();
        probe[n++] = acc.getValue(nanovdb::Coord(o[0] + 1, o[1] + 2, o[2] + 3));
    }
    for (; n < 3; ++n) probe[n] = 0.f;
    probe[3] = acc.getValue(nanovdb::Coord(box.min()[0] + 100, box.min()[1] + 100, box.min()[2] + 100));
    printf("{\"mode\": \"info\", \"active\": %llu, \"min\": [%d, %d, %d], \"max\": [%d, %d, %d], \"leaves\": %u, \"probe\": [%.9g, %.9g, %.9g, %.9g]}\n",
           (unsigned long long) grid->activeVoxelCount(), box.min()[0], box.min()[1], box.min()[2], box.max()[0], box.max()[1], box.max()[2],
           (unsigned) grid->tree().nodeCount(0), probe[0], probe[1], probe[2], probe[3]);
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 2) return fail("usage: nvdb_tool write in.nbv out.nvdb [name] | dump in.nvdb out.bin [grid name]");
    try {
        int rc = 2;
        if (!strcmp(argv[1], "write")) rc = modeWrite(argc, argv);
        else if (!strcmp(argv[1], "dump")) rc = modeDump(argc, argv);
        else if (!strcmp(argv[1], "info")) rc = modeInfo(argc, argv);
        if (rc == 2) return fail("usage: nvdb_tool write in.nbv out.nvdb [name] | dump in.nvdb out.bin [grid name]");
        return rc;
    } catch (const std::exception &e) {
        return fail(e.what());
    }
}
