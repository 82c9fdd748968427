/* bvh.cpp — SAH BVH build, restating the reference's builder (src/utils/bvh.cpp:100-380) so that
 * scenes loaded here get the same tree the reference would traverse:
 *   >= 32 primitives: split the largest axis by 16 centroid bins, SAH cost
 *        2*T + (nL*SA(L) + nR*SA(R)) / SA(node) against the leaf cost n, stable partition;
 *   <  32 primitives (or no binned split found): exact sweep over all three axes with std::sort;
 *   children at idx+1 and idx+2*left_count in a 2n array, then compaction (bvh.cpp:357-377).
 * The reference runs this under TBB; the order inside a partition is then scheduling dependent.
 * Here (and in the oracle's serial TBB stand-in) ranges are processed in ascending order, which
 * makes the partition stable and the tree deterministic. fp32 throughout, no FMA contraction
 * (-ffp-contract=off in build.py).
 * Like the reference's TBB tasks, subtrees are built side by side: a subtree owns disjoint ranges of the node array,
 * of the index array and of the scratch array, so a large right child is handed to another thread while this one goes
 * on with the left child; the result does not depend on the schedule (same bytes with 1 or 16 threads). */
#include "plugins.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace nori {

namespace {

enum { SERIAL_THRESHOLD = 32, BIN_COUNT = 16, TRAVERSAL_COST = 1, INTERSECTION_COST = 1 };

struct Node {                /* BVH::BVHNode, include/nori/bvh.h:127-164 */
    uint32_t word0 = 0, word1 = 0;
    BoundingBox3f bbox;
    bool used = false;
    void setLeaf(uint32_t start, uint32_t size) { word0 = 1u | (size << 1); word1 = start; used = true; }
    void setInner(uint32_t axis, uint32_t right) { word0 = (axis << 1); word1 = right; used = true; }
    bool isLeaf() const { return word0 & 1u; }
};

struct Builder {
    const std::vector<Shape *> &shapes;
    std::vector<uint32_t> shapeOffset;
    std::vector<Node> nodes;
    std::vector<uint32_t> indices, temp;
    /* centroids and boxes are pure functions of the primitive: computed once */
    std::vector<Point3f> centroid;
    std::vector<BoundingBox3f> box;
    std::atomic<int> spare_threads{0};         /* worker threads that may still be started */
    std::atomic<int> forks{0};
    enum { PARALLEL_THRESHOLD = 8192 };        /* primitives below which a subtree is not worth a thread */

    explicit Builder(const std::vector<Shape *> &s) : shapes(s) {}

    void serial(uint32_t node_idx, uint32_t *start, uint32_t *end, uint32_t *tmp) {       /* bvh.cpp:236-305 */
        Node &node = nodes[node_idx];
        const uint32_t size = (uint32_t) (end - start);
        float best_cost = (float) INTERSECTION_COST * size;
        int64_t best_index = -1, best_axis = -1;
        float *left_areas = (float *) tmp;
        for (int axis = 0; axis < 3; ++axis) {
            std::sort(start, end, [&](uint32_t f1, uint32_t f2) { return centroid[f1][axis] < centroid[f2][axis]; });
            BoundingBox3f bbox;
            for (uint32_t i = 0; i < size; ++i) { bbox.expandBy(box[start[i]]); left_areas[i] = bbox.getSurfaceArea(); }
            if (axis == 0) node.bbox = bbox;
            bbox.reset();
            const float tri_factor = INTERSECTION_COST / node.bbox.getSurfaceArea();
            for (uint32_t i = size - 1; i >= 1; --i) {
                bbox.expandBy(box[start[i]]);
                const float left_area = left_areas[i - 1], right_area = bbox.getSurfaceArea();
                const uint32_t prims_left = i, prims_right = size - i;
                const float sah_cost = 2.0f * TRAVERSAL_COST + tri_factor * (prims_left * left_area + prims_right * right_area);
                if (sah_cost < best_cost) { best_cost = sah_cost; best_index = i; best_axis = axis; }
            }
        }
        if (best_index == -1) {
            node.setLeaf((uint32_t) (start - indices.data()), size);
            return;
        }
        std::sort(start, end, [&](uint32_t f1, uint32_t f2) { return centroid[f1][best_axis] < centroid[f2][best_axis]; });
        const uint32_t left_count = (uint32_t) best_index;
        const uint32_t left = node_idx + 1, right = node_idx + 2 * left_count;
        node.setInner((uint32_t) best_axis, right);
        /* the two halves own disjoint node, index and scratch ranges: a large one goes to another thread. (On a mesh
         * where the 16-bin split of a big node beats nothing — C3: a 1.19 M-triangle node — the whole subtree is built
         * by this sweep, as in the reference, and that is where the build time is.) */
        const uint32_t right_count = size - left_count;
        if (left_count >= PARALLEL_THRESHOLD && right_count >= PARALLEL_THRESHOLD) {
            if (spare_threads.fetch_sub(1) > 0) {
                forks.fetch_add(1);
                uint32_t *const r_start = start + left_count, *const r_tmp = tmp + left_count;
                std::thread worker([this, right, r_start, end, r_tmp]() { serial(right, r_start, end, r_tmp); });
                serial(left, start, start + left_count, tmp);
                worker.join();
                spare_threads.fetch_add(1);
                return;
            }
            spare_threads.fetch_add(1);
        }
        serial(left, start, start + left_count, tmp);
        serial(right, start + left_count, end, tmp + left_count);
    }

    void binned(uint32_t node_idx, uint32_t *start, uint32_t *end, uint32_t *tmp) {       /* bvh.cpp:100-233 */
        for (;;) {
            const uint32_t size = (uint32_t) (end - start);
            Node &node = nodes[node_idx];
            node.used = true;
            if (size < SERIAL_THRESHOLD) { serial(node_idx, start, end, tmp); return; }
            const int axis = node.bbox.getLargestAxis();
            const float mn = node.bbox.min[axis], mx = node.bbox.max[axis], inv_bin_size = BIN_COUNT / (mx - mn);
            uint32_t counts[BIN_COUNT];
            BoundingBox3f bbox[BIN_COUNT];
            memset(counts, 0, sizeof(counts));
            for (uint32_t i = 0; i < size; ++i) {
                const uint32_t f = start[i];
                const int index = std::min(std::max((int) ((centroid[f][axis] - mn) * inv_bin_size), 0), (int) BIN_COUNT - 1);
                counts[index]++;
                bbox[index].expandBy(box[f]);
            }
            BoundingBox3f bbox_left[BIN_COUNT];
            bbox_left[0] = bbox[0];
            for (int i = 1; i < BIN_COUNT; ++i) { counts[i] += counts[i - 1]; bbox_left[i] = BoundingBox3f::merge(bbox_left[i - 1], bbox[i]); }
            BoundingBox3f bbox_right = bbox[BIN_COUNT - 1], best_bbox_right;
            int64_t best_index = -1;
            float best_cost = (float) INTERSECTION_COST * size;
            const float tri_factor = (float) INTERSECTION_COST / node.bbox.getSurfaceArea();
            for (int i = BIN_COUNT - 2; i >= 0; --i) {
                const uint32_t prims_left = counts[i], prims_right = size - counts[i];
                const float sah_cost = 2.0f * TRAVERSAL_COST +
                    tri_factor * (prims_left * bbox_left[i].getSurfaceArea() + prims_right * bbox_right.getSurfaceArea());
                if (sah_cost < best_cost) { best_cost = sah_cost; best_index = i; best_bbox_right = bbox_right; }
                bbox_right = BoundingBox3f::merge(bbox_right, bbox[i]);
            }
            if (best_index == -1) { if (getenv("NB_HOST_TIMING") && size > 4096) fprintf(stderr, "binned split failed at %u primitives -> sweep\n", size); serial(node_idx, start, end, tmp); return; }
            const uint32_t left_count = counts[best_index];
            const uint32_t left = node_idx + 1, right = node_idx + 2 * left_count;
            nodes[left].bbox = bbox_left[best_index];
            nodes[right].bbox = best_bbox_right;
            node.setInner((uint32_t) axis, right);
            uint32_t idx_l = 0, idx_r = left_count;
            for (uint32_t i = 0; i < size; ++i) {                        /* stable partition on the unclamped bin */
                const uint32_t f = start[i];
                const int index = (int) ((centroid[f][axis] - mn) * inv_bin_size);
                if (index <= best_index) tmp[idx_l++] = f; else tmp[idx_r++] = f;
            }
            memcpy(start, tmp, size * sizeof(uint32_t));
            uint32_t *const r_start = start + left_count, *const r_end = end, *const r_tmp = tmp + left_count;
            const bool fork = (uint32_t) (r_end - r_start) >= PARALLEL_THRESHOLD && left_count >= PARALLEL_THRESHOLD &&
                              spare_threads.fetch_sub(1) > 0;
            if (fork) {
                forks.fetch_add(1);
                std::thread worker([this, right, r_start, r_end, r_tmp]() { binned(right, r_start, r_end, r_tmp); });
                binned(left, start, start + left_count, tmp);
                worker.join();
                spare_threads.fetch_add(1);
                return;
            }
            if ((uint32_t) (r_end - r_start) >= PARALLEL_THRESHOLD && left_count >= PARALLEL_THRESHOLD) spare_threads.fetch_add(1);   /* undo the failed claim */
            binned(right, r_start, r_end, r_tmp);
            node_idx = left;                                             /* continue with the left subtree */
            end = start + left_count;
        }
    }
};

} // namespace

void buildBVH(const std::vector<Shape *> &shapes, std::vector<nb_bvh_node> &out_nodes, std::vector<uint32_t> &out_indices) {
    Builder b(shapes);
    b.shapeOffset.push_back(0);
    BoundingBox3f root;
    for (Shape *s : shapes) {                                            /* BVH::addShape, bvh.cpp:308-317 */
        b.shapeOffset.push_back(b.shapeOffset.back() + s->getPrimitiveCount());
        root.expandBy(s->getBoundingBox());
    }
    const uint32_t size = b.shapeOffset.back();
    out_nodes.clear(); out_indices.clear();
    if (size == 0) return;
    b.centroid.resize(size); b.box.resize(size);
    for (size_t s = 0; s < shapes.size(); ++s)
        for (uint32_t i = 0; i < shapes[s]->getPrimitiveCount(); ++i) {
            b.centroid[b.shapeOffset[s] + i] = shapes[s]->getCentroid(i);
            b.box[b.shapeOffset[s] + i] = shapes[s]->getBoundingBox(i);
        }
    b.nodes.assign(2 * (size_t) size, Node());
    b.nodes[0].bbox = root;
    b.indices.resize(size); b.temp.resize(size);
    for (uint32_t i = 0; i < size; ++i) b.indices[i] = i;
    {
        const char *env = getenv("NB_BVH_THREADS");
        int threads = env ? atoi(env) : (int) std::thread::hardware_concurrency();
        if (threads < 1) threads = 1;
        b.spare_threads = threads - 1;
        const auto t0 = std::chrono::steady_clock::now();
        b.binned(0, b.indices.data(), b.indices.data() + size, b.temp.data());
        if (getenv("NB_HOST_TIMING"))
            fprintf(stderr, "buildBVH: %u primitives, %d threads, %d forks, %.3f s\n", size, threads, b.forks.load(),
                    std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }

    /* compaction (bvh.cpp:357-377): drop unused slots, keep order, re-point right children */
    std::vector<uint32_t> remap(b.nodes.size(), 0);
    uint32_t n = 0;
    for (size_t j = 0; j < b.nodes.size(); ++j) if (b.nodes[j].used) remap[j] = n++;
    out_nodes.resize(n);
    for (size_t j = 0; j < b.nodes.size(); ++j) {
        if (!b.nodes[j].used) continue;
        const Node &src = b.nodes[j];
        nb_bvh_node &dst = out_nodes[remap[j]];
        dst.word0 = src.word0;
        dst.word1 = src.isLeaf() ? src.word1 : remap[src.word1];
        for (int k = 0; k < 3; ++k) { dst.bmin[k] = src.bbox.min[k]; dst.bmax[k] = src.bbox.max[k]; }
    }
    out_indices = b.indices;
}

} // namespace nori
